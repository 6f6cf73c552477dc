"""The NCCL / copy-engine rotation on real hardware: an N-rank run (one process per GPU under
torch.distributed.run) must reproduce the loss curve of the same schedule walked on ONE GPU
(same buckets, same per-bucket order keys; only the Hogwild interleaving differs) and stay within
1 % of the single-thread oracle walking that block order.  Needs >= 2 GPUs: skipped on the
1-GPU test box, run with `gpurun --gpus 2`.  Replaces the thread slicing of
J/opt/Optimizer.java:59-63."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, tmp_path, rot=None, V=20000, E=160000, dim=32, epochs=6, layout="contiguous"):
    out = tmp_path / f"dist_{world}_{rot}_{layout}.json"
    env = dict(os.environ)
    env["GE_TEST_LAYOUT"] = layout
    if rot:
        env["GE_DIST_ROT"] = rot
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_worker.py"), str(out), str(V), str(E), str(dim), str(epochs)]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    with open(out) as f:
        return json.load(f)


def _worlds():
    import torch
    n = torch.cuda.device_count()
    only = os.environ.get("GE_TEST_WORLDS")        # e.g. "8": one world size per (expensive) 8-GPU call
    return [w for w in (2, 4, 8) if w <= n and (not only or str(w) in only.split(","))]


@pytest.mark.parametrize("rot,layout", [("ce", "contiguous"), ("nccl", "contiguous"), ("ce", "cyclic")])
def test_n_rank_run_matches_one_rank_schedule(tmp_path, rot, layout):
    worlds = _worlds()
    if not worlds:
        pytest.skip("needs at least 2 GPUs")
    for world in worlds:
        r = _run(world, tmp_path, rot, layout=layout)
        assert r["layout"] == layout
        multi, single, rep = np.array(r["hist"]), np.array(r["single"]), np.array(r["single_repeat"])
        noise = np.abs(rep - single) / single
        rel = np.abs(multi - single) / single
        print(f"world={world} rotation={r['rotation']} rel={np.round(rel, 5)} run-to-run noise={np.round(noise, 5)} "
              f"cos={r['cosine_vs_single']:.4f} (repeat {r['cosine_single_repeat']:.4f})")
        assert r["finite"]
        if rot == "ce":
            assert "copy engine" in r["rotation"], r["rotation"]      # no silent fallback to NCCL
        if rot == "nccl" and world > 1:
            # P = 1 buckets: other order keys than the 1-rank P = 2 walk -> statistical agreement only
            assert rel.max() < 0.01, rel
        else:
            # identical buckets and order keys: what is left is Hogwild's run-to-run noise, which on this small
            # graph reaches 0.5 % in the first epoch (two 1-rank runs differ by as much) and 0.1 % later
            # (seen over the round's runs: epoch 1 up to 1.0 %, epoch 2 up to 0.4 %, later epochs <= 0.05 %)
            assert rel[0] < 0.02 and rel[1] < 0.008 and rel[2:].max() < max(2e-3, 3 * noise[2:].max()), (rel, noise)
        assert r["cosine_vs_single"] > min(0.97, r["cosine_single_repeat"] - 0.02)
