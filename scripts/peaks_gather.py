#!/usr/bin/env python
"""Gather / vector-reduction peak probes (SURVEY.md section 8d "gather roofline").

Runs de_probe_gather (deepearth_b200/csrc/misc.cu) over working sets that live in L2 (32 MiB) and in
HBM (4 GiB), for 4/8/16-byte entries, loads and red.add.f32 vectors, with and without the x-pair
pattern, and prints one JSON document with the op rates (G ops/s) and the sector bytes they imply.
Needs a GPU: run under gpurun; the committed copy of its output is profiles/peaks_gather_r01.json.
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from deepearth_b200 import _lib  # noqa: E402


def probe(lib, table, n_entries, entry_bytes, mode, n_threads, per_thread, pair, sink, reps=5):
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = (ctypes.c_void_p(table.data_ptr()), n_entries, entry_bytes, mode, n_threads, per_thread, pair,
            ctypes.c_void_p(sink.data_ptr()), st)
    for _ in range(2):
        _lib.check(lib.de_probe_gather(*args), "probe")
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.de_probe_gather(*args), "probe")
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return n_threads * per_thread / (best * 1e-3) / 1e9, best


def main():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    sink = torch.zeros(4, device=dev)
    out = {"gpu": torch.cuda.get_device_name(0), "unit": "G ops/s", "rows": []}
    n_threads = 148 * 2048 * 4
    for ws_name, ws_bytes in (("L2_2MiB", 2 << 20), ("L2_32MiB", 32 << 20), ("L2_64MiB", 64 << 20), ("HBM_4GiB", 4 << 30)):
        buf = torch.zeros(ws_bytes // 4, dtype=torch.float32, device=dev)
        for entry_bytes in (4, 8, 16):
            n_entries = ws_bytes // entry_bytes
            for mode, mname in ((0, "load"), (1, "red.add.f32")):
                for pair in (0, 1, 2, 4, 8):
                    rate, ms = probe(lib, buf, n_entries, entry_bytes, mode, n_threads, 64, pair, sink)
                    out["rows"].append({"working_set": ws_name, "entry_bytes": entry_bytes, "op": mname,
                                        "pattern": {0: "random", 1: "x-pair (consecutive instr)", 2: "2 adjacent lanes", 4: "4 adjacent lanes", 8: "8 adjacent lanes"}[pair], "gops": round(rate, 2), "ms": round(ms, 4),
                                        "useful_GBps": round(rate * entry_bytes, 1)})
                    print(out["rows"][-1], flush=True)
        del buf
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "peaks_gather.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
