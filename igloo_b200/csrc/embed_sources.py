"""Build step: embeds the device headers NVRTC needs into jit_sources.inc
(name, raw string literal pairs).  System includes are rewritten to the quoted form so
they resolve against the headers fp_jit.cpp supplies."""
import re
import sys

out, files = sys.argv[1], sys.argv[2:]
with open(out, "w") as f:
    for path in files:
        name = path.split("/")[-1]
        src = open(path).read()
        src = re.sub(r'#include\s*<(stdint|stddef)\.h>', r'#include "\1.h"', src)
        assert ')IGLSRC"' not in src
        f.write('{"%s", R"IGLSRC(%s)IGLSRC"},\n' % (name, src))
