"""Per-role clock64 timeline of CTA (0,0) of gemm_tc_kernel (needs a -DB200IR_TIMELINE build: see csrc/build.sh).
Usage: B200IR_LIB=tools/ab/lib_tl.so python tools/gemm_timeline.py k1 l1b [first_tile n_tiles]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.argv_saved = sys.argv[:]
sys.argv = sys.argv[:1]
import torch  # noqa: E402
import b200ir  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_ops  # noqa: E402

NAMES = {1: "P.empty", 2: "P.issued", 10: "M.accempty", 11: "M.full", 12: "M.commit", 20: "S.full", 21: "S.done",
         22: "S.stempty", 29: "E.top", 30: "E.stfull", 31: "E.accfull", 32: "E.bulkwait", 34: "E.math", 35: "E.arrive",
         36: "E.stored"}


def main():
    a = sys.argv_saved[1:]
    op, lv = a[0], a[1]
    t0, nt = (int(a[2]), int(a[3])) if len(a) > 3 else (20, 3)
    lib = ctypes.CDLL(b200ir._lib.LIB_PATH)
    lib.b200ir_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
    fn, _ = bench_ops.make(op, lv)
    for i in range(3):
        fn(i)
    lib.b200ir_debug_timeline(None, 0)   # sync + reset
    fn(0)
    buf = (ctypes.c_ulonglong * (2 * 16384))()
    n = lib.b200ir_debug_timeline(buf, 16384)
    ev = sorted(((buf[2 * i + 1], buf[2 * i] >> 32, buf[2 * i] & 0xffffffff) for i in range(n) if buf[2 * i + 1]))
    n = len(ev)
    base = ev[0][0]
    print(f"{n} events, span {ev[-1][0] - base} cycles, last tile id {max(e[2] for e in ev) // 8}")
    first = min(e[2] for e in ev) // 8
    for clk, tag, ident in ev:
        tile = ident // 8 - first
        if t0 <= tile < t0 + nt:
            print(f"{clk - base:9d}  tile {tile:3d}.{ident % 8}  {NAMES.get(tag, tag)}")


if __name__ == "__main__":
    main()
