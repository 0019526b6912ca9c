"""Issue-rate micro-benchmarks (os1_b200/csrc/ubench.cu -> libos1_ubench.so): what bench.py holds the Hamming kernels to."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libos1_ubench.so")


def measure(device=0):
    lib = C.CDLL(LIB_PATH)
    lib.os1_ubench_issue_rates.argtypes = [C.c_int, C.POINTER(C.c_double)]
    out = (C.c_double * 11)()
    rc = lib.os1_ubench_issue_rates(device, out)
    if rc != 0:
        raise RuntimeError("os1_ubench_issue_rates failed")
    v = list(out)
    return {"popc_per_clk_per_sm": v[0], "popc_per_s": v[1], "lop3_per_clk_per_sm": v[2], "lop3_per_s": v[3],
            "dist256_per_clk_per_sm": v[4], "dist256_per_s": v[5], "scan_pairs_per_clk_per_sm": v[6],
            "hamming_pairs_per_s_ceiling": v[7], "sm_mhz_during_scan_loop": v[8],
            "scan_pairs_per_s_3csa_5popc": v[9], "scan_pairs_per_s_2csa_6popc": v[10],
            "how": "os1_b200/csrc/ubench.cu at 64 warps per SM; the ceiling is the matchers' inner loop exactly (row in registers, candidates "
                   "from a resident shared-memory tile, 8 XOR + 4 carry-save adders + 4 POPC + key and best/second update) with no tile "
                   "refills, barriers, tails or merge"}
