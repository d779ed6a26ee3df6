"""Data-parallel path on real GPUs (needs >= 2 B200 on the box; skipped otherwise): N shards with in-kernel
synchronised BatchNorm and the fused reduce-scatter / Adam / all-gather == one process on the global batch (SURVEY.md
section 8c multi-GPU oracle).  Cases: the small shape and the shape bench.py times (1,024 wide, 2 blocks, 8,192 rows per
rank), world 2 / 4 / 8, CUDA-graph replay on and off, per-layer optimiser ranges on and off, replicas constructed from
different seeds.  bench.py --gpus N repeats the check on the timed path itself (``dp_check``)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [  # world, case, use_graph, fuse_optim, divergent_init
    (2, "small", "1", "1", "0"), (2, "bench", "1", "1", "0"), (2, "bench", "1", "0", "0"), (2, "small", "0", "1", "0"),
    (2, "small", "1", "1", "1"), (4, "bench", "1", "1", "0"), (8, "bench", "1", "1", "0"),
    (2, "vit", "1", "1", "0"), (2, "vit", "0", "1", "0"), (4, "vit", "1", "1", "0"),      # the transformer lifter
]


@pytest.mark.parametrize("world,case,use_graph,fuse,divergent", CASES)
def test_data_parallel_matches_single_process(world, case, use_graph, fuse, divergent):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, DP_CASE=case, DP_USE_GRAPH=use_graph, LIFTER_FUSE_OPTIM=fuse, DP_DIVERGENT_INIT=divergent,
               LIFTER_DP_TIMEOUT_S="30")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "dp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
