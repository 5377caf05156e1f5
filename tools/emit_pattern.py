#!/usr/bin/env python
"""Emit orbslam3lib_b200/csrc/orb_pattern.h from the reference's rBRIEF table.

Run once in the build container (needs /root/reference); the emitted header is committed.
The table is the OpenCV ORB `bit_pattern_31_` data (reference copy:
app/src/main/cpp/src/ORBextractor_old.cc:151-409).  It is data both sides must share for the
descriptors to be comparable, so it is extracted mechanically instead of being retyped.
"""
import hashlib
import re
import sys

REF = "/root/reference/app/src/main/cpp/src/ORBextractor_old.cc"
OUT = "orbslam3lib_b200/csrc/orb_pattern.h"


def main():
    src = open(REF).read()
    m = re.search(r"bit_pattern_31_\[256\*4\]\s*=\s*\{(.*?)\};", src, re.S)
    body = re.sub(r"/\*.*?\*/", "", m.group(1), flags=re.S)
    body = re.sub(r"//.*", "", body)
    vals = [int(v) for v in re.findall(r"-?\d+", body)]
    assert len(vals) == 1024, len(vals)
    digest = hashlib.sha256(bytes([(v + 256) % 256 for v in vals])).hexdigest()
    lines = ["    " + ",".join("%d" % v for v in vals[i:i + 16]) + "," for i in range(0, 1024, 16)]
    hdr = (
        "// ORB rotated-BRIEF sampling pattern: 256 test pairs (x0,y0,x1,y1), signed 8-bit offsets in [-13,13].\n"
        "// This is the learned rBRIEF table that OpenCV ships in modules/features2d/src/orb.cpp and that the\n"
        "// reference repeats at app/src/main/cpp/src/ORBextractor_old.cc:151-409 (and dsp/src/orbslam_dsp_rb.h:44-302).\n"
        "// It is interoperability DATA, not logic: descriptors are only comparable between systems that use these\n"
        "// exact offsets.  Emitted by tools/emit_pattern.py; sha256 of the int8 bytes is checked in tests/test_pattern.py.\n"
        "#pragma once\n"
        "static const signed char ORBFE_BIT_PATTERN_31[1024] = {\n" + "\n".join(lines) + "\n};\n"
    )
    open(OUT, "w").write(hdr)
    print(digest)


if __name__ == "__main__":
    sys.exit(main())
