#!/usr/bin/env python
"""tcgen05 known-answer tests + issue-rate / hop-latency microbenchmarks (ll_tc05_probe, csrc/ll_tc05_probe.cu).
Prints one JSON object per case; run on a B200:  python scripts/tc05_probe.py > gpurun_out/tc05_probe.jsonl"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = C.CDLL(os.path.join(ROOT, "lunar_lander_b200", "libll_gpu.so"))
lib.ll_tc05_probe.argtypes = [C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_double)]
lib.ll_last_error.restype = C.c_char_p
lib.ll_last_error.argtypes = [C.c_void_p]


def probe(test, m=128, n=32, k=16, a_mn=0, b_mn=0, swap=0, iters=0, n_acc=1, busy_warps=0, busy_iters=0, mufu_every=0,
          group_threads=128, all_sms=0, ts=0, lane_off=0):
    args = (C.c_int32 * 16)(test, m, n, k, a_mn, b_mn, swap, iters, n_acc, busy_warps, busy_iters, mufu_every, group_threads, all_sms, ts, lane_off)
    out = np.zeros((128, n), dtype=np.float32)
    res = (C.c_double * 16)()
    rc = lib.ll_tc05_probe(0, args, out.ctypes.data_as(C.POINTER(C.c_float)), res)
    if rc != 0:
        return {"rc": rc, "err": (lib.ll_last_error(None) or b"").decode()}, out
    return {"rc": 0, "res": [float(x) for x in res[:5]]}, out


def ref(m_rows, n, k):
    a = np.array([[((r * 7 + kk * 3) % 13) - 6 for kk in range(k)] for r in range(128)], dtype=np.float64)
    b = np.array([[((r * 5 + kk * 11) % 9) - 4 for kk in range(k)] for r in range(n)], dtype=np.float64)
    return a @ b.T


def main():
    only = sys.argv[1:]  # e.g. "kat", "rate", "overlap", "hop"
    # ---- known answers: every operand-major combination, with both LBO / SBO conventions
    if not only or "kat" in only:
        for (m, n, k) in ((128, 32, 16), (128, 32, 128), (128, 64, 32), (128, 16, 32), (64, 32, 32)):
            for a_mn in (0, 1):
                for b_mn in (0, 1):
                    for swap in (0, 1):
                        r, d = probe(0, m=m, n=n, k=k, a_mn=a_mn, b_mn=b_mn, swap=swap)
                        rec = {"case": "kat", "m": m, "n": n, "k": k, "a_mn": a_mn, "b_mn": b_mn, "swap": swap, **r}
                        if r["rc"] == 0 and m == 64:
                            # where did the 64 rows land? lane -> matching reference row (or -1)
                            want = ref(128, n, k)
                            lanes = []
                            for lane in range(128):
                                hit = [row for row in range(64) if np.array_equal(d[lane].astype(np.float64), want[row])]
                                lanes.append(hit[0] if hit else -1)
                            rec["lane_to_row"] = lanes
                        print(json.dumps(rec), flush=True)
    # ---- A operand from TMEM (tcgen05.st by the row's own thread), M = 128 and the two interleaved M = 64 tiles
    if not only or "ts" in only:
        for (m, n, k, lane_off) in ((128, 32, 16, 0), (128, 32, 128, 0), (128, 128, 32, 0), (128, 16, 32, 0), (64, 32, 128, 0), (64, 32, 128, 16),
                                    (64, 128, 32, 0), (64, 128, 32, 16), (64, 8, 32, 16), (64, 16, 16, 0)):
            if n % 16:
                continue
            r, d = probe(0, m=m, n=n, k=k, ts=1, lane_off=lane_off)
            print(json.dumps({"case": "kat_ts", "m": m, "n": n, "k": k, "lane_off": lane_off, **r}), flush=True)
        for m in (64, 128):
            for n in (8, 16, 32, 64, 128):
                if m == 128 and n % 16:
                    continue
                for n_acc in (1, 2):
                    r, _ = probe(1, m=m, n=n, iters=32768, n_acc=n_acc, all_sms=1, ts=1)
                    print(json.dumps({"case": "rate_ts", "m": m, "n": n, "n_acc": n_acc, **r}), flush=True)
        for m in (64, 128):
            for n in (8, 16, 32, 64, 128):
                if m == 128 and n % 16:
                    continue
                for per_hop in (1, 8, 32, 128):
                    r, _ = probe(2, m=m, n=n, iters=1000, n_acc=per_hop, group_threads=128, ts=1)
                    print(json.dumps({"case": "hop_ts", "m": m, "n": n, "mma_per_hop": per_hop, **r}), flush=True)
    # ---- issue rate, tensor pipe alone (n_acc rotating accumulators, a power of two)
    if not only or "rate" in only:
        for m in (64, 128):
            for n in (8, 16, 32, 64, 128, 256):
                if m == 128 and n % 16:
                    continue
                for n_acc in (1, 4):
                    if n_acc * n > 512:
                        continue
                    r, _ = probe(1, m=m, n=n, iters=32768, n_acc=n_acc, all_sms=1)
                    print(json.dumps({"case": "rate", "m": m, "n": n, "n_acc": n_acc, **r}), flush=True)
    # ---- the epilogue-like stream alone, then next to the tensor pipe (does tcgen05.mma steal issue slots?)
    if not only or "overlap" in only:
        for busy_warps in (4, 8, 12):
            for mufu_every in (0, 4, 1):
                base, _ = probe(1, m=128, n=32, iters=0, busy_warps=busy_warps, busy_iters=20000, mufu_every=mufu_every, all_sms=1)
                both, _ = probe(1, m=128, n=32, iters=1 << 20, n_acc=4, busy_warps=busy_warps, busy_iters=20000, mufu_every=mufu_every, all_sms=1)
                print(json.dumps({"case": "overlap", "busy_warps": busy_warps, "mufu_every": mufu_every, "busy_alone": base, "with_mma": both}), flush=True)
    # ---- hop latency: n_acc = tcgen05.mma per hop
    if not only or "hop" in only:
        for g in (32, 128):
            for n in (16, 32, 64):
                for per_hop in (1, 6, 24):
                    r, _ = probe(2, m=128, n=n, iters=2000, n_acc=per_hop, group_threads=g)
                    print(json.dumps({"case": "hop", "group_threads": g, "n": n, "mma_per_hop": per_hop, **r}), flush=True)


if __name__ == "__main__":
    sys.exit(main())
