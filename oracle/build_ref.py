#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — builds oracle/_ref/libptref.so, the *reference itself* compiled for the CPU.

What is compiled (nothing here is product code, and no reference source is committed):
  * Kernel/PathTracer_FullKernel.cl + _header.cl — the reference megakernel, compiled as C++17 through
    oracle/shim/ocl_as_cxx.h after a mechanical rewrite of a *generated copy*:
        (float4)(a,b,c,d)   ->  float4{a,b,c,d}         OpenCL vector literals (brace-init keeps the
        (float2)(..) (int2)(..)                          left-to-right evaluation order)
        #define FLOAT4 (float4) -> #define FLOAT4 make_float4   (same for FLOAT2/INT4/RGBACOLOR)
    and with oracle/shim/X2inv.cl supplied for the reference's empty Kernel/X2inv.cl.
    Every arithmetic statement is the reference's own.
  * Controleur/PathTracer_BVH.cpp (+ the 4 headers it includes) — the reference BVH builder, compiled
    from a generated copy with one rewrite:  `ALIGN(16) typedef struct`  ->  `typedef struct alignas(16)`
    (MSVC __declspec(align) has no g++ spelling that binds to an anonymous struct behind a typedef),
    and an empty stand-in for <CL/cl.h>.
  * oracle/ref_glue.cpp — C entry points around both.

The generated copies live in oracle/_ref/gen/ only while compiling and are deleted afterwards
(--keep-gen keeps them). oracle/_ref/ is git-ignored but travels to the GPU box.

Usage:  python oracle/build_ref.py [--reference /root/reference] [--keep-gen]
Exit code 0 and prints the .so path; exit code 3 if the reference tree is absent (GPU box): the
prebuilt .so is then used as-is if present.
"""
import argparse
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")

VEC_TYPES = ("float2", "float4", "float8", "int2", "int4", "uint2", "RGBAColor")
_LIT = re.compile(r"\((%s)\)\s*\(" % "|".join(VEC_TYPES))
_DEF = re.compile(r"^(#define\s+\w+\s+)\((%s)\)\s*$" % "|".join(VEC_TYPES), re.M)


def rewrite_vector_literals(src: str) -> str:
    """(T)(a, b, ...) -> T{a, b, ...} with parenthesis matching; `#define X (T)` -> `#define X make_T`."""
    src = _DEF.sub(lambda m: "%smake_%s" % (m.group(1), m.group(2)), src)
    out = []
    i = 0
    while True:
        m = _LIT.search(src, i)
        if not m:
            out.append(src[i:])
            break
        out.append(src[i:m.start()])
        j = m.end()  # just past the opening '(' of the argument list
        depth = 1
        while depth:
            c = src[j]
            if c == "(":
                depth += 1
            elif c == ")":
                depth -= 1
            j += 1
        inner = rewrite_vector_literals(src[m.end():j - 1])
        if m.group(1) == "float8":
            # only ever appears as an argument of the (compiled-out) ASSERT macros, where braces would
            # not protect the commas from the preprocessor
            out.append("make_float8(%s)" % inner)
        else:
            out.append("%s{%s}" % (m.group(1), inner))
        i = j
    return "".join(out)


def rewrite_align(src: str) -> str:
    return re.sub(r"ALIGN\((\d+)\)\s+typedef\s+struct", r"typedef struct alignas(\1)", src)


def read_latin1(path):
    with open(path, "r", encoding="latin-1") as f:
        return f.read()


def write_latin1(path, text):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w", encoding="latin-1") as f:
        f.write(text)


def build(reference: str, keep_gen: bool = False, verbose: bool = True) -> str:
    kdir = os.path.join(reference, "Kernel")
    cdir = os.path.join(reference, "Controleur")
    if not os.path.isdir(kdir):
        raise FileNotFoundError(reference)
    if os.path.isdir(GEN):
        shutil.rmtree(GEN)
    # --- generated kernel copy -----------------------------------------------------------
    for name in ("PathTracer_FullKernel.cl", "PathTracer_FullKernel_header.cl"):
        write_latin1(os.path.join(GEN, "kernel", name), rewrite_vector_literals(read_latin1(os.path.join(kdir, name))))
    shutil.copy(os.path.join(HERE, "shim", "X2inv.cl"), os.path.join(GEN, "kernel", "X2inv.cl"))
    # --- generated host copy (BVH builder + the headers it pulls) ------------------------
    for name in ("PathTracer_BVH.cpp", "PathTracer_BVH.h", "PathTracer_Utils.h", "PathTracer_Structs.h",
                 "PathTracer_PreProc.h"):
        write_latin1(os.path.join(GEN, "host", name), rewrite_align(read_latin1(os.path.join(cdir, name))))

    so = os.path.join(OUT, "libptref.so")
    common = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fwrapv", "-fPIC", "-fopenmp",
              "-Wno-narrowing", "-Wno-unused", "-w"]
    objs = []
    cmds = [
        # reference BVH builder, unmodified apart from the ALIGN rewrite
        common + ["-c", os.path.join(GEN, "host", "PathTracer_BVH.cpp"), "-I", os.path.join(HERE, "shim"),
                  "-include", "climits", "-include", "cstring", "-include", "stdexcept",
                  "-D__declspec(x)=", "-o", os.path.join(GEN, "bvh.o")],
        # glue + kernel (three sampler variants inside)
        common + ["-c", os.path.join(HERE, "ref_glue.cpp"), "-I", os.path.join(HERE, "shim"),
                  "-I", os.path.join(GEN, "kernel"), "-I", os.path.join(GEN, "host"),
                  "-D__declspec(x)=", "-o", os.path.join(GEN, "glue.o")],
    ]
    for c in cmds:
        if verbose:
            print(" ".join(c), flush=True)
        subprocess.check_call(c)
        objs.append(c[-1])
    link = ["g++", "-shared", "-fopenmp", "-o", so] + objs
    if verbose:
        print(" ".join(link), flush=True)
    subprocess.check_call(link)
    if not keep_gen:
        shutil.rmtree(GEN)
    return so


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--keep-gen", action="store_true")
    a = ap.parse_args()
    if not os.path.isdir(os.path.join(a.reference, "Kernel")):
        print("reference tree not present at %s; not rebuilding oracle/_ref" % a.reference)
        sys.exit(3)
    print(build(a.reference, a.keep_gen))


if __name__ == "__main__":
    main()
