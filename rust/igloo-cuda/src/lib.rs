//! igloo-cuda: GPU (NVIDIA B200, sm_100a) implementations of the engine's Filter, Projection
//! and HashJoin physical operators behind the engine's own `ExecutionPlan` trait.
//! UNBUILT in this repository's environment (no Rust toolchain) — see README.md.
pub mod expr;
pub mod device;
pub mod ffi;
pub mod ops;

pub use ffi::GpuContext;
pub use ops::{GpuFilterExec, GpuHashJoinExec, GpuProjectionExec};
