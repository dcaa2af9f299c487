"""Per-kernel CUDA-event durations UNDER LOAD: S sequences on T worker threads, the profile switched on in every context
(two event records per launch: the numbers are device-side durations, not throughput)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ldso_b200 import capi  # noqa: E402

S, T = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "16:8").split(":"))
steps, warm = 20, 5
seeds = [1001 + i for i in range(S)]
data = bench.render_sequences(seeds, bench.PRIME + warm + steps + 1)
import torch  # noqa: E402
gpu = capi.load()
pinned = {s: [torch.from_numpy(bench.to_raw8(f)).pin_memory() for f in data[s][0]] for s in seeds}
job = bench.Job(gpu, seeds, data, threads=T)
ptr = lambda i, k: pinned[seeds[i]][k].data_ptr()
job.run(ptr, bench.PRIME + warm, bench.RAW8, True)
for c in job.ctxs:
    c.profile_enable(True)
job.run(ptr, steps, bench.RAW8, True)
tot = {}
for c in job.ctxs:
    for (name, which, key, unit) in bench.PROF:
        ms, n, units = c.profile_get(which)
        a = tot.setdefault(name, [0.0, 0])
        a[0] += ms; a[1] += n
print("S %d threads %d: per-kernel durations under load" % (S, T))
for name, (ms, n) in tot.items():
    if n:
        print("  %-28s launches %5d  avg %8.1f us  per frame %7.1f us" % (name, n, 1e3 * ms / n, 1e3 * ms / (S * steps)))
job.close()
