"""Writes the named member-function DEFINITIONS of a reference source file, verbatim, to a scratch file OUTSIDE the
repository (oracle/Makefile target `ref_frame` compiles that file and deletes it).  TEST INFRASTRUCTURE ONLY.

Why: /root/reference/src/Frame.cc cannot be compiled as a whole in this image (its constructors, IMU and BoW members pull
in Eigen, Sophus, g2o, boost and the real OpenCV), but the functions on the hot path can: each one is cut from its
signature line to the closing brace in column 0 and compiled unmodified against the stand-in `Frame` of
ref_shim_matcher/prelude.h.  Nothing of the reference is written under /root/repo.

usage: extract_ref_functions.py <source.cc> <out.inc> <signature prefix> [<signature prefix> ...]"""
import sys


def main():
    src, out, sigs = sys.argv[1], sys.argv[2], sys.argv[3:]
    lines = open(src).read().split("\n")
    chunks = []
    for sig in sigs:
        starts = [i for i, l in enumerate(lines) if l.lstrip().startswith(sig)]
        if len(starts) != 1:
            sys.exit("extract_ref_functions: %d definitions start with %r in %s" % (len(starts), sig, src))
        a = starts[0]
        indent = lines[a][:len(lines[a]) - len(lines[a].lstrip())]   # the closing brace sits at the signature's indentation
        b = next(i for i in range(a + 1, len(lines)) if lines[i].rstrip() == indent + "}")
        chunks.append("#line %d \"%s\"\n" % (a + 1, src) + "\n".join(lines[a:b + 1]) + "\n")
    open(out, "w").write("\n".join(chunks))


if __name__ == "__main__":
    main()
