"""TEST INFRASTRUCTURE: writes the reference's src/Frame.cc with the body of Frame::ComputeStereoMatches() replaced by
oracle/gpu_shim/ComputeStereoMatches.body.inc (step 3 of INTEGRATION.md) to the path given as argv[2]. The output is a
build intermediate in a temporary directory (oracle/Makefile deletes it after compiling); nothing of the reference's
source is kept in the repository. Usage: make_frame_gpu.py /root/reference/src/Frame.cc /tmp/xyz/Frame.cc"""
import os
import sys

src, dst = sys.argv[1], sys.argv[2]
text = open(src).read()
head = "void Frame::ComputeStereoMatches()"
at = text.index(head)
assert text.count(head) == 1
open_brace = text.index("{", at)
depth, i = 0, open_brace
while True:
    c = text[i]
    if c == "{":
        depth += 1
    elif c == "}":
        depth -= 1
        if depth == 0:
            break
    i += 1
body = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ComputeStereoMatches.body.inc")).read()
open(dst, "w").write(text[:open_brace] + body.rstrip("\n") + text[i + 1:])
