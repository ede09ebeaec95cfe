//! `Arc<dyn PhysicalExpr>` -> postfix bytecode (include/igloo_expr.h).
//!
//! The planner hands the operators expressions that DataFusion has already type-coerced
//! (physical_planner.rs:60-64, 86-90, 107-116), so both sides of a binary operator agree;
//! igl_expr_compile re-checks.  Anything outside the subset returns `NotImplemented`, and
//! the planner keeps the CPU operator for that node.  igloo_b200/expr.py is the same
//! lowering in Python, which the tests exercise.
use arrow::datatypes::DataType;
use datafusion::common::ScalarValue;
use datafusion::error::{DataFusionError, Result as DataFusionResult};
use datafusion::logical_expr::Operator;
use datafusion::physical_expr::expressions::{
    BinaryExpr, CastExpr, Column, IsNotNullExpr, IsNullExpr, Literal, NegativeExpr, NotExpr,
};
use datafusion::physical_expr::PhysicalExpr;
use std::sync::Arc;

const COL: u8 = 0x01;
const LIT_BOOL: u8 = 0x02;
const LIT_I32: u8 = 0x03;
const LIT_I64: u8 = 0x04;
const LIT_F64: u8 = 0x05;
const LIT_NULL: u8 = 0x06;
const NEG: u8 = 0x25;
const NOT: u8 = 0x32;
const IS_NULL: u8 = 0x33;
const IS_NOT_NULL: u8 = 0x34;
const CAST: u8 = 0x40;

fn type_code(t: &DataType) -> DataFusionResult<u8> {
    Ok(match t {
        DataType::Boolean => 1,
        DataType::Int32 => 2,
        DataType::Int64 => 3,
        DataType::Float64 => 4,
        DataType::Utf8 => 5,
        other => return Err(DataFusionError::NotImplemented(format!("igloo-cuda: type {other:?}"))),
    })
}

fn binary_opcode(op: &Operator) -> DataFusionResult<u8> {
    Ok(match op {
        Operator::Eq => 0x10,
        Operator::NotEq => 0x11,
        Operator::Lt => 0x12,
        Operator::LtEq => 0x13,
        Operator::Gt => 0x14,
        Operator::GtEq => 0x15,
        Operator::Plus => 0x20,
        Operator::Minus => 0x21,
        Operator::Multiply => 0x22,
        Operator::Divide => 0x23,
        Operator::Modulo => 0x24,
        Operator::And => 0x30,
        Operator::Or => 0x31,
        other => return Err(DataFusionError::NotImplemented(format!("igloo-cuda: operator {other:?}"))),
    })
}

/// Appends the postfix encoding of `expr` to `out`.
pub fn encode(expr: &Arc<dyn PhysicalExpr>, out: &mut Vec<u8>) -> DataFusionResult<()> {
    let any = expr.as_any();
    if let Some(c) = any.downcast_ref::<Column>() {
        out.push(COL);
        out.extend_from_slice(&(c.index() as u16).to_le_bytes());
    } else if let Some(l) = any.downcast_ref::<Literal>() {
        match l.value() {
            ScalarValue::Boolean(Some(v)) => { out.push(LIT_BOOL); out.push(*v as u8); }
            ScalarValue::Int32(Some(v)) => { out.push(LIT_I32); out.extend_from_slice(&v.to_le_bytes()); }
            ScalarValue::Int64(Some(v)) => { out.push(LIT_I64); out.extend_from_slice(&v.to_le_bytes()); }
            ScalarValue::Float64(Some(v)) => { out.push(LIT_F64); out.extend_from_slice(&v.to_bits().to_le_bytes()); }
            null if null.is_null() => { out.push(LIT_NULL); out.push(type_code(&null.data_type())?); }
            other => return Err(DataFusionError::NotImplemented(format!("igloo-cuda: literal {other:?}"))),
        }
    } else if let Some(b) = any.downcast_ref::<BinaryExpr>() {
        encode(b.left(), out)?;
        encode(b.right(), out)?;
        out.push(binary_opcode(b.op())?);
    } else if let Some(n) = any.downcast_ref::<NotExpr>() {
        encode(n.arg(), out)?;
        out.push(NOT);
    } else if let Some(n) = any.downcast_ref::<NegativeExpr>() {
        encode(n.arg(), out)?;
        out.push(NEG);
    } else if let Some(n) = any.downcast_ref::<IsNullExpr>() {
        encode(n.arg(), out)?;
        out.push(IS_NULL);
    } else if let Some(n) = any.downcast_ref::<IsNotNullExpr>() {
        encode(n.arg(), out)?;
        out.push(IS_NOT_NULL);
    } else if let Some(c) = any.downcast_ref::<CastExpr>() {
        encode(c.expr(), out)?;
        out.push(CAST);
        out.push(type_code(c.cast_type())?);
    } else {
        return Err(DataFusionError::NotImplemented(format!("igloo-cuda: expression {expr:?}")));
    }
    Ok(())
}

pub fn to_bytecode(expr: &Arc<dyn PhysicalExpr>) -> DataFusionResult<Vec<u8>> {
    let mut v = Vec::with_capacity(16);
    encode(expr, &mut v)?;
    Ok(v)
}
