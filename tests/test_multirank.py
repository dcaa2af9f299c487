"""N > 1 path of bench.py on CPU: two gloo ranks, each stepping ITS OWN sequences (independent sequences, no data-path
collective — SURVEY.md section 8e) through the C++ orchestrator over the oracle; the only exchange is the max-over-ranks
of the timed region and the gather of the per-rank logs."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import bench
    from ldso_b200 import build, capi
    from tests import util
    bench.CALIB, bench.PRIME, bench.STRIDE = util.CALIB["small"], 24, 1           # a small image keeps the CPU test short
    dist.init_process_group("gloo", rank=rank, world_size=world)
    api = capi.Api(build.ORACLE_LIB, "orc_")
    seeds = bench.rank_seeds(rank, 2)
    data = bench.render_sequences(seeds, bench.PRIME + 1 + 3 + 1)
    job = bench.Job(api, seeds, data)
    frames = {s: [np.ascontiguousarray(f, np.float32) for f in data[s][0]] for s in seeds}
    ptr = lambda i, k: frames[seeds[i]][k].ctypes.data
    job.prime(ptr, bench.PRIME)
    times, _, _, _, first = bench.timed_region(job, ptr, False, 3, 1)
    st = job.steady_state(first, first + 3)
    local = sum(times) + 10.0 * rank                       # make the slowest rank unambiguous
    (t_max,), per_rank = bench.reduce_over_ranks([local], dict(rank=rank, seeds=seeds, ate=st["ate_rmse_m_max"], lost=st["lost"]),
                                                 world, dist, torch, "cpu")
    cen = job.sys[0].centres()[1][-1].tolist()
    out.put((rank, local, t_max, per_rank, cen))
    job.close()
    dist.destroy_process_group()


def test_two_ranks_shard_by_sequence_and_take_max_time():
    from ldso_b200 import build
    build.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=300) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, l0, m0, pr0, c0), (r1, l1, m1, pr1, c1) = res
    assert m0 == m1 == max(l0, l1)                         # every rank sees the slowest rank's time
    assert pr0 == pr1 and [p["rank"] for p in pr0] == [0, 1]
    assert pr0[0]["seeds"] == [1001, 1002] and pr0[1]["seeds"] == [1003, 1004]     # disjoint sequences per rank
    assert not np.allclose(c0, c1)                         # different seeds -> different sequences -> different poses
    assert all(p["lost"] == 0 and p["ate"] < 0.05 for p in pr0)                    # both ranks really ran their sequences
