"""N > 1 path of bench.py on CPU: two gloo ranks, each stepping ITS OWN sequence (independent sequences, no data-path
collective — SURVEY.md section 8e) through the oracle; the only exchange is the max-over-ranks of the timed region."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
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
    import bench
    from ldso_b200 import build, capi
    dist.init_process_group("gloo", rank=rank, world_size=world)
    api = capi.Api(build.ORACLE_FAST_LIB, "orc_")
    seq = bench.Sequence(api, bench.rank_seed(rank), 2)
    seq.prime()
    times = bench.run_sequence_timed(seq, 2, 0, use_dev=False)[0]
    local = sum(times) + 10.0 * rank                       # make the slowest rank unambiguous
    dist.barrier()
    t_max, = bench.max_over_ranks([local], world, dist, torch, "cpu")
    pose = seq.results[-1][3]
    out.put((rank, local, t_max, bench.aggregate_fps(world, 2, t_max), pose.tolist(), seq.gt_error()))
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
    (r0, l0, m0, v0, p0, e0), (r1, l1, m1, v1, p1, e1) = res
    assert m0 == m1 == max(l0, l1)                         # every rank sees the slowest rank's time
    assert v0 == v1 == 2 * 2 / m0                          # whole-job frames/s over both ranks
    assert not np.allclose(p0, p1)                         # different seeds -> different sequences -> different poses
    assert e0 < 0.02 and e1 < 0.02                         # both ranks really tracked their sequence
