#!/usr/bin/env python
"""tools/embed_headers.py OUT HEADER... -- the kernel headers as C++ raw string literals
({"name", R"PINTJIT(text)PINTJIT"},), included by pint_b200/csrc/jit.cu: the library carries the
templates it specialises at run time, so NVRTC needs no source tree."""
import os
import sys

out, headers = sys.argv[1], sys.argv[2:]
chunks = []
for h in headers:
    text = open(h).read()
    assert ")PINTJIT\"" not in text
    # a raw string literal may be long, but keep every piece under the 64 KB some front ends insist on
    pieces = [text[i:i + 60000] for i in range(0, len(text), 60000)]
    lit = "\n".join('R"PINTJIT(%s)PINTJIT"' % p for p in pieces)
    chunks.append('{"%s",\n%s},\n' % (os.path.basename(h), lit))
new = "".join(chunks)
if not os.path.exists(out) or open(out).read() != new:
    os.makedirs(os.path.dirname(out), exist_ok=True)
    open(out, "w").write(new)
