"""Aggregate frames/s on ONE GPU against the number of sequences stepped together (bench.py's Job / timed_region),
for tuning: LDSO_LM_CLUSTER (CTAs per tracker cluster) and the number of worker threads."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ldso_b200 import capi  # noqa: E402

# usage: seq_scaling.py S_MAX            -> S = 1, 2, 4, ... S_MAX with S and S/2 worker threads
#        seq_scaling.py S1,S2,...        -> exactly these S, one worker thread per sequence (LDSO_WAIT_YIELD=1 to oversubscribe)
#        seq_scaling.py S1:T1,S2:T2,...  -> S sequences on T worker threads (a worker begins the frames of all its sequences
#                                           before it finishes the first: include/ldso_system.hpp BatchRunner)
PAIRS = None
LIST = None
if len(sys.argv) > 1 and ":" in sys.argv[1]:
    PAIRS = [tuple(int(v) for v in x.split(":")) for x in sys.argv[1].split(",")]
    LIST = sorted({p[0] for p in PAIRS})
elif len(sys.argv) > 1 and "," in sys.argv[1]:
    LIST = [int(x) for x in sys.argv[1].split(",")]
S_MAX = max(LIST) if LIST else (int(sys.argv[1]) if len(sys.argv) > 1 else 8)
steps, warm = 20, 5
seeds = [1001 + i for i in range(S_MAX)]
data = bench.render_sequences(seeds, bench.PRIME + warm + steps + 1)
if os.environ.get("LDSO_WAIT_BLOCK", "0") != "0":   # the primary context is created with blocking synchronisation
    import ctypes
    ctypes.CDLL("libcudart.so.12").cudaSetDeviceFlags(4)
import torch  # noqa: E402
gpu = capi.load()
DEV = os.environ.get("SEQ_DEVICE_IMAGES", "0") != "0"      # images resident in HBM (bench.py's `value`) instead of page-locked host memory (e2e)
pinned = {s: [torch.from_numpy(f).cuda() if DEV else torch.from_numpy(bench.to_raw8(f)).pin_memory() for f in data[s][0]] for s in seeds}
KIND = 1 if DEV else bench.RAW8   # e2e input: the 8-bit sensor image in page-locked host memory
for S in (LIST or [1, 2, 4, 8, 16][: [1, 2, 4, 8, 16].index(S_MAX) + 1]):
    for threads in ([p[1] for p in PAIRS if p[0] == S] if PAIRS else [S] if LIST else sorted({S, max(1, S // 2)}, reverse=True)):
        sd = seeds[:S]
        job = bench.Job(gpu, sd, data, threads=threads)
        ptr = lambda i, k: pinned[sd[i]][k].data_ptr()
        job.run(ptr, bench.PRIME, KIND, True)
        t, launches, h2d, d2h, first = bench.timed_region(job, ptr, KIND, steps, warm)
        st = job.steady_state(first, first + steps)
        print("S %2d threads %2d: %.0f frames/s, %.3f ms/step, %.3f ms/frame/seq, kf %d" %
              (S, threads, S * steps / sum(t), 1e3 * sum(t) / steps, 1e3 * sum(t) / steps / S, st["keyframes_in_timed_region"]), flush=True)
        job.close()
