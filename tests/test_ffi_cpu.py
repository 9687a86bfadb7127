"""CPU: the C-ABI library loads and exports every symbol include/psob.h declares; host-side
sharding / record protocol, including a world_size-2 gloo run."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "psob.h")).read()
    return sorted(set(re.findall(r"\b(psob_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from psob import _ffi
    if not os.path.exists(_ffi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    l = _ffi.lib()
    syms = _header_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(l, s), f"{s} declared in include/psob.h but not exported"
    assert sorted(_ffi.EXPORTS) == syms
    assert l.psob_version() == 200


def test_no_cpu_fallback():
    from psob import _ffi, engine
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_ffi.PsobError):
        engine.require_cuda()
    from src.utils import util_dcgan
    g = util_dcgan.Generator(10, 1, 16)
    with pytest.raises(_ffi.PsobError):
        g(torch.zeros(2, 10, 1, 1))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gan-discovery-pso_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, os.path.join(dp, f)


def test_shard_ranges_cover_swarm():
    from psob import dist as pd
    for total, world in ((1048576, 8), (65536, 3), (10, 4)):
        spans = [pd.shard_range(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == total
        for (a, ca), (b, _) in zip(spans[:-1], spans[1:]):
            assert a + ca == b


def test_record_protocol_matches_global_scan(golden):
    from oracle import port
    from psob import dist as pd
    rng = np.random.RandomState(3)
    P, D, world = 64, 8, 4
    val = (rng.random_sample(P).astype(np.float32) + 0.1)
    val[40] = val[7] = val.min()                      # tie across shards -> lowest global index
    pos = rng.randn(P, D).astype(np.float32)
    recs = []
    for r in range(world):
        first, cnt = pd.shard_range(P, r, world)
        recs.append(pd.make_record(val[first:first + cnt], pos[first:first + cnt], first))
    win = pd.select_record(np.stack(recs))
    ref_val, ref_idx = port.gbest_scan(val, np.inf)
    assert int(np.stack(recs)[win].view(np.uint32)[0]) == ref_idx == 7
    assert np.stack(recs)[win][2] == ref_val
    np.testing.assert_array_equal(np.stack(recs)[win][4:], pos[ref_idx])


_GLOO_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.join(sys.argv[1], "gan-discovery-pso_b200"))
from psob import dist as pd
rank, world, _ = pd.init_from_env("gloo")
rng = np.random.RandomState(11)
P, D = 50, 12
val = rng.random_sample(P).astype(np.float32) + 0.1
val[33] = val[2] = val.min()
pos = rng.randn(P, D).astype(np.float32)
first, cnt = pd.shard_range(P, rank, world)
rec = torch.from_numpy(pd.make_record(val[first:first+cnt], pos[first:first+cnt], first))
recs = torch.zeros(world, 4 + D)
dist.all_gather_into_tensor(recs.view(-1), rec)
win = pd.select_record(recs.numpy())
idx = int(recs.numpy()[win].view(np.uint32)[0])
assert idx == 2, idx
assert np.array_equal(recs.numpy()[win][4:], pos[2])
dist.barrier()
print("rank", rank, "ok")
'''


def test_gbest_exchange_world2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29611")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29611", str(script), ROOT],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2
