"""Stateful fuzz of DeviceTable: random sequences of add / delete / compact / snapshot round trips, each followed by
filtered searches that are compared with a NumPy model of the table (float64 oracle on the surviving rows under the
same WHERE clause).  usage: python tools/fuzz_table.py [sequences] [seed]"""
import asyncio
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dab_b200 import search_service as ss  # noqa: E402
from oracle import search_ref  # noqa: E402

DEV = "cuda:0"
STATUS = ["completed", "processing", "failed"]


def one_sequence(rng, seq):
    dim = int(rng.choice([32, 64, 384]))
    dtype = torch.bfloat16 if rng.random() < 0.5 else torch.float32
    t = ss.DeviceTable(6000, dim, dtype, DEV)
    model = dict(pk=np.zeros(0, np.int64), doc=np.zeros(0, np.int64), bot=np.zeros(0, np.int64), status=np.zeros(0, object),
                 emb=np.zeros((0, dim), np.float32), alive=np.zeros(0, bool))
    next_pk = 100
    for step in range(int(rng.integers(4, 10))):
        op = rng.choice(["add", "add", "delete", "compact", "snapshot"]) if len(model["pk"]) else "add"
        if op == "add" and len(model["pk"]) < 5000:
            n = int(rng.integers(1, 700))
            emb = rng.standard_normal((n, dim)).astype(np.float32)
            pks = np.arange(next_pk, next_pk + n, dtype=np.int64)
            next_pk += n + int(rng.integers(0, 5))
            docs = rng.integers(0, 40, n)
            bots = rng.integers(0, 3, n)
            status = np.asarray([STATUS[i] for i in rng.integers(0, 3, n)], dtype=object)
            t.add(pks.tolist(), emb, docs.tolist(), [f"t{p}" for p in pks], None, document__wiki__bot=bots.tolist(),
                  document__processing__status=status.tolist())
            stored = torch.from_numpy(emb).to(dtype).float().numpy()
            model = dict(pk=np.concatenate([model["pk"], pks]), doc=np.concatenate([model["doc"], docs]),
                         bot=np.concatenate([model["bot"], bots]), status=np.concatenate([model["status"], status]),
                         emb=np.concatenate([model["emb"], stored]), alive=np.concatenate([model["alive"], np.ones(n, bool)]))
        elif op == "delete":
            live = np.flatnonzero(model["alive"])
            if len(live) > 10:
                gone = rng.choice(live, size=int(rng.integers(1, max(2, len(live) // 4))), replace=False)
                assert t.delete(model["pk"][gone].tolist() + [999_999_999]) == len(gone)
                model["alive"][gone] = False
        elif op == "compact":
            removed = t.compact()
            assert removed == int((~model["alive"]).sum())
            keep = model["alive"]
            model = {k: v[keep] for k, v in model.items()}
        elif op == "snapshot":
            with tempfile.TemporaryDirectory() as d:
                path = os.path.join(d, "snap.npz")
                t.save(path)
                t.close()
                t = ss.DeviceTable.load(path, capacity=6000, device=DEV)
        # --- queries against the model
        if not model["alive"].any():
            continue
        q = rng.standard_normal(dim).astype(np.float32)
        kind = int(rng.integers(0, 5))
        qs = ss.ChunkQuerySet(t)
        mask = model["alive"].copy()
        if kind == 1:
            b = int(rng.integers(0, 3))
            qs, mask = qs.filter(document__wiki__bot=b), mask & (model["bot"] == b)
        elif kind == 2:
            lim = int(rng.integers(0, 40))
            qs, mask = qs.filter(document__id__lt=lim).exclude(document__processing__status="failed"), mask & (model["doc"] < lim) & (model["status"] != "failed")
        elif kind == 3:
            some = rng.choice(model["pk"], size=min(len(model["pk"]), int(rng.integers(1, 50))), replace=False)
            qs, mask = qs.filter(pk__in=some.tolist()), mask & np.isin(model["pk"], some)
        elif kind == 4:
            qs, mask = qs.filter(document__processing__status__in=["completed", "processing"]), mask & (model["status"] != "failed")
        n = int(rng.choice([1, 5, 10, 250]))
        hits = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=n))
        S = search_ref.exact_scores(model["emb"], q[None])[0]
        want, wsc = search_ref.rank_topk(S, n, mask)
        got_rows = np.asarray([int(np.flatnonzero(model["pk"] == h.pk)[0]) for h in hits], dtype=np.int64)
        search_ref.check_topk_parity(S, got_rows, np.asarray([1.0 - h.distance for h in hits]), n, mask)
        assert len(hits) == len(want), (len(hits), len(want))
        assert all(h.text == f"t{h.pk}" and h.document_id == model["doc"][r] for h, r in zip(hits, got_rows))
        assert qs.count() == int(mask.sum())
    t.close()
    return step


def main():
    seqs = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(seed)
    for s in range(seqs):
        try:
            one_sequence(rng, s)
        except Exception:
            print(f"sequence {s} (seed {seed}) failed", flush=True)
            raise
    print(f"{seqs} sequences passed (seed {seed})")


if __name__ == "__main__":
    main()
