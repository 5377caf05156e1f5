#!/usr/bin/env python
"""Compile the REFERENCE'S OWN CPU extractor glue into oracle/_ref/libext_old.so.  Test infrastructure.

    python oracle/ref_ext/make_ref_ext.py [--keep-generated DIR]        (also: make -C oracle ref_ext)

Reads, where they lie under /root/reference (never copied into the repo: the generated translation unit lives in a
temporary directory that is removed after the compile; only the .so lands in the git-ignored oracle/_ref/):

  app/src/main/cpp/src/ORBextractor_old.cc          the extractor (EXT_OLD)
  app/src/main/cpp/include/ORBextractor_old.h       its class declaration

and applies exactly these edits, each needed because the file as committed upstream does not compile anywhere
(the shipped build runs this code on the DSP instead):

  header  drop the two LynxHardwareAcceleration includes (Android / FastRPC / GLES headers) and forward-declare
          AHardwareBuffer, which only the dropped second operator() uses
  :477    remove the stray token `ExtractorNode` after `{` in computeOrientation
  :900-1077   drop ComputeKeyPointsOld (dead code: its call at :1111 is commented out)
  :1104   `LynxHardwareAccelerator::GetInstance().ComputePyramid(image, nlevels, mvInvScaleFactor, mvImagePyramid);`
          (a method that no longer exists, LynxHardwareAccelerator.h) -> the line the authors left commented out
          right below it, :1105 `ComputePyramid(image);`
  :1193-1327  drop the AHardwareBuffer operator() (DSP path; its body refers to removed members)
  :1331-1356  uncomment ORBextractor::ComputePyramid (the original ORB-SLAM3 body, commented out verbatim)

Everything else — IC_Angle :78-105, computeOrbDescriptor :109-148, bit_pattern_31_ :151-409, the constructor
:411-471, computeOrientation :473-480, DivideNode :482-538, compareNodes :540-555, DistributeOctTree :557-781,
ComputeKeyPointsOctTree :783-898, computeDescriptors :1079-1086, operator() :1088-1191 — is compiled as written,
against oracle/ref_ext/cv_shim.h (OpenCV containers + the five OpenCV algorithms forwarded to the cv2-pinned
primitives of liborb_oracle.so) and oracle/ref_ext/shim/android/log.h (log lines -> nothing).

Compiler: g++ / libstdc++ (std::sort = introsort).  The shipped Android build uses libc++, whose std::sort orders
compareNodes-equivalent elements differently; see DESIGN.md §2.
"""
import os
import re
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE = os.path.dirname(HERE)
REF = os.environ.get("ORBFE_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "app/src/main/cpp/src/ORBextractor_old.cc")
HDR = os.path.join(REF, "app/src/main/cpp/include/ORBextractor_old.h")
OUT = os.path.join(ORACLE, "_ref", "libext_old.so")


def expect(cond, what):
    if not cond:
        raise SystemExit("make_ref_ext: the reference is not the file this recipe was written for: " + what)


def generate(dst):
    lines = open(SRC, encoding="utf-8").read().split("\n")
    L = lambda n: lines[n - 1]    # 1-based, as cited

    expect(len(lines) >= 1358 and "ORBextractor::operator()( InputArray _image" in L(1088), "operator() at :1088")
    expect(L(477).strip() == "{ExtractorNode", ":477 stray token")
    expect("ORBextractor::ComputeKeyPointsOld" in L(900) and L(1077).strip() == "}", "ComputeKeyPointsOld :900-1077")
    expect("LynxHardwareAccelerator::GetInstance().ComputePyramid" in L(1104), ":1104 removed method")
    expect(L(1105).strip() == "//        ComputePyramid(image);", ":1105 commented call")
    expect("AHardwareBuffer* _image" in L(1195) and L(1327).strip() == "}", "AHardwareBuffer operator() :1195-1327")
    expect(L(1331).strip() == "//    void ORBextractor::ComputePyramid(cv::Mat image)" and L(1356).strip() == "//    }",
           "commented ComputePyramid :1331-1356")

    out = []
    for n in range(1, len(lines) + 1):
        s = L(n)
        if n == 477:
            s = s.replace("{ExtractorNode", "{")
        elif 900 <= n <= 1077 or 1193 <= n <= 1327:
            s = ""    # keep the line numbers of the original for diagnostics
        elif n == 1104:
            s = ""
        elif n == 1105:
            s = s.replace("//", "  ", 1)
        elif 1331 <= n <= 1356:
            s = re.sub(r"^//", "  ", s)
        out.append(s)
    open(os.path.join(dst, "ORBextractor_old_gen.cc"), "w", encoding="utf-8").write("\n".join(out))

    h = open(HDR, encoding="utf-8").read()
    for inc in ('#include "LynxHardwareAcceleration/LynxHardwareAccelerator.h"',
                '#include "LynxHardwareAcceleration/FrameAHB.h"'):
        expect(inc in h, inc)
        h = h.replace(inc, "")
    h = h.replace("namespace ORB_SLAM3\n{", "struct AHardwareBuffer;\nnamespace ORB_SLAM3\n{", 1)
    # the .cc includes "ORBextractor.h"; the class that matches it is the one declared in ORBextractor_old.h
    open(os.path.join(dst, "ORBextractor.h"), "w", encoding="utf-8").write(h)


def main():
    if not (os.path.exists(SRC) and os.path.exists(HDR)):
        print("oracle/_ref: %s not present (GPU box): libext_old.so skipped" % REF)
        return 0
    keep = None
    if "--keep-generated" in sys.argv:
        keep = sys.argv[sys.argv.index("--keep-generated") + 1]
    subprocess.check_call(["make", "-C", ORACLE, "-s"])    # liborb_oracle.so: the forwarded primitives
    tmp = keep or tempfile.mkdtemp(prefix="ref_ext_")
    os.makedirs(tmp, exist_ok=True)
    try:
        generate(tmp)
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        cxx = os.environ.get("CXX", "g++")
        cmd = [cxx, "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-std=c++17", "-w", "-shared",
               "-I", tmp, "-I", os.path.join(HERE, "shim"), "-I", HERE,
               os.path.join(tmp, "ORBextractor_old_gen.cc"), os.path.join(HERE, "ref_ext_api.cpp"),
               "-L", os.path.join(ORACLE, "_build"), "-lorb_oracle", "-Wl,-rpath,$ORIGIN/../_build", "-o", OUT]
        subprocess.check_call(cmd)
    finally:
        if not keep:
            shutil.rmtree(tmp, ignore_errors=True)
    print("built", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
