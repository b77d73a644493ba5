#!/usr/bin/env python
"""Host-to-device copy bandwidth of this box, measured with bare cudaMemcpyAsync (no engine): each GPU
alone, then 2 / 4 / 8 GPUs at once, with one and two copies in flight per GPU, one large copy against
64 MiB chunks, and write-combined against default pinned memory.  Answers whether the end-to-end scaling
of bench.py (H2D inside the timed region) is limited by the platform.

    python tools_h2d_matrix.py [--mib 1024] > profiles/r02_h2d_matrix.json
"""
import argparse
import ctypes
import glob
import json
import os
import threading

import torch


def cudart():
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            return ctypes.CDLL(name)
        except OSError:
            pass
    import nvidia.cuda_runtime as m   # the wheel torch depends on
    return ctypes.CDLL(glob.glob(os.path.join(os.path.dirname(m.__file__), "lib", "libcudart.so*"))[0])


RT = cudart()
RT.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
RT.cudaFreeHost.argtypes = [ctypes.c_void_p]
RT.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
H2D = 1
PORTABLE, WRITE_COMBINED = 1, 4


def host_alloc(nbytes, wc):
    p = ctypes.c_void_p()
    rc = RT.cudaHostAlloc(ctypes.byref(p), nbytes, PORTABLE | (WRITE_COMBINED if wc else 0))
    assert rc == 0, rc
    ctypes.memset(p, 0x61, nbytes)
    return p


def run(gpus, nbytes, in_flight, chunk, wc, reps=4):
    """GB/s per GPU (and summed) when every GPU of `gpus` copies `in_flight` buffers of nbytes at once."""
    host = {g: [host_alloc(nbytes, wc) for _ in range(in_flight)] for g in gpus}
    dev = {g: [torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % g) for _ in range(in_flight)] for g in gpus}
    streams = {g: [torch.cuda.Stream(device=g) for _ in range(in_flight)] for g in gpus}
    result = {}
    start = threading.Barrier(len(gpus))

    def one(g):
        torch.cuda.set_device(g)
        best = 0.0
        for _ in range(reps):
            torch.cuda.synchronize(g)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.wait()
            e0.record(torch.cuda.current_stream(g))
            for k in range(in_flight):
                st = streams[g][k]
                st.wait_event(e0)
                for off in range(0, nbytes, chunk):
                    n = min(chunk, nbytes - off)
                    rc = RT.cudaMemcpyAsync(dev[g][k].data_ptr() + off, host[g][k].value + off, n, H2D, ctypes.c_void_p(st.cuda_stream))
                    assert rc == 0, rc
            for k in range(in_flight):
                torch.cuda.current_stream(g).wait_stream(streams[g][k])
            e1.record(torch.cuda.current_stream(g))
            torch.cuda.synchronize(g)
            best = max(best, in_flight * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9)
        result[g] = best

    threads = [threading.Thread(target=one, args=(g,)) for g in gpus]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for g in gpus:
        for p in host[g]:
            RT.cudaFreeHost(p)
    return {"gpus": list(gpus), "copies_in_flight_per_gpu": in_flight, "chunk_mib": chunk >> 20, "write_combined": wc,
            "gbs_per_gpu": [round(result[g], 2) for g in gpus], "gbs_total": round(sum(result.values()), 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    args = ap.parse_args()
    nbytes = args.mib << 20
    n = torch.cuda.device_count()
    rows = []
    for g in range(n):
        rows.append(run([g], nbytes, 1, nbytes, False))
    for k in (2, 4, 8):
        if k > n:
            break
        for in_flight in (1, 2):
            for chunk in (nbytes, 64 << 20):
                rows.append(run(list(range(k)), nbytes, in_flight, chunk, False))
        rows.append(run(list(range(k)), nbytes, 1, nbytes, True))
    rows.append(run([0], nbytes, 2, nbytes, False))
    rows.append(run([0], nbytes, 1, nbytes, True))
    print(json.dumps({"what": "pinned host -> device cudaMemcpyAsync, best of 4, CUDA events per GPU, all listed GPUs copying at once",
                      "bytes_per_copy": nbytes, "cpu_count": os.cpu_count(), "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
