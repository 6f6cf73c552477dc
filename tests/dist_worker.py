"""Worker of tests/test_gpu_dist_multi.py: one rank of an N-GPU block-rotation run under
torch.distributed.run.  Rank 0 also walks the same schedule alone (world = 1, rotation = pointer
shift) and writes both loss curves."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from graph_embeddings_b200 import BCA, Configuration, Opt, synth
    from graph_embeddings_b200.dist import DistTrainer

    out_path, V, E, dim, epochs = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    method = sys.argv[6] if len(sys.argv) > 6 else "glove"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = synth.rmat(V, E, 4)
    cfg = Configuration(method=method, dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=31, device=local)
    idb = [DistTrainer.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(idb, src=0)
    layout = os.environ.get("GE_TEST_LAYOUT", "contiguous")
    tr = DistTrainer(cfg, rank, world, 8, idb[0], layout=layout)
    tr.build(g)
    tr.prepare()
    hist = [tr.epoch() for _ in range(epochs)]
    emb = tr.extract_local()
    ids = tr.local_ids()
    assert np.array_equal(ids, tr.owned_ids(g.V))
    parts = [None] * world if rank == 0 else None
    dist.gather_object((ids, emb), parts, dst=0)
    res = {"world": world, "hist": hist, "rotation": tr.stats()["rotation"], "layout": layout}
    dist.barrier()
    if rank == 0:
        full = np.zeros((g.V, dim), np.float32)
        seen = np.zeros(g.V, np.int32)
        for ids_r, e in parts:
            full[ids_r] = e
            seen[ids_r] += 1
        assert (seen == 1).all()                   # the ranks' rows partition the vertices
        single = []
        embs = []
        for rep in range(2):                       # twice: the second run measures Hogwild run-to-run noise
            t1 = DistTrainer(cfg, 0, 1, 8, None, layout=layout)
            t1.build(g)
            t1.prepare()
            single.append([t1.epoch() for _ in range(epochs)])
            e1 = np.zeros((g.V, dim), np.float32)
            e1[t1.local_ids()] = t1.extract_local()
            embs.append(e1)
            del t1
        a, b = full.astype(np.float64).ravel(), embs[0].astype(np.float64).ravel()
        res.update(single=single[0], single_repeat=single[1],
                   cosine_vs_single=float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b))),
                   cosine_single_repeat=float(embs[0].astype(np.float64).ravel() @ embs[1].astype(np.float64).ravel()
                                              / (np.linalg.norm(embs[0]) * np.linalg.norm(embs[1]))),
                   finite=bool(np.isfinite(full).all()))
        with open(out_path, "w") as f:
            json.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
