"""CPU, world_size 2, gloo: the sweep's head partition and its single all_gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cram_b200.sweep import assemble_gathered, gather_ie, heads_of_rank


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_ie(l, h):
    return float(l * 100 + h) + 0.25


def _worker(rank, world, port, L, H, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = heads_of_rank(H, rank, world)
    width = (H + world - 1) // world
    local = torch.full((L, width), float("nan"))
    for l in range(L):
        for i, h in enumerate(mine):
            local[l, i] = _fake_ie(l, h)
    full = gather_ie(local, L, H, world)
    q.put((rank, full))
    dist.barrier()
    dist.destroy_process_group()


def test_partition_covers_every_head_once():
    for H in (4, 28, 32, 40, 64):
        for N in (1, 2, 3, 4, 8):
            seen = sorted(h for r in range(N) for h in heads_of_rank(H, r, N))
            assert seen == list(range(H))
            sizes = [len(heads_of_rank(H, r, N)) for r in range(N)]
            assert max(sizes) - min(sizes) <= 1


def test_assemble_is_inverse_of_partition():
    L, H, N = 3, 28, 8
    width = (H + N - 1) // N
    parts = []
    for r in range(N):
        t = torch.full((L, width), float("nan"))
        for i, h in enumerate(heads_of_rank(H, r, N)):
            t[:, i] = torch.arange(L) * 100.0 + h
        parts.append(t)
    full = assemble_gathered(parts, L, H, N)
    assert torch.equal(full, torch.arange(L)[:, None] * 100.0 + torch.arange(H)[None, :])


def test_two_rank_gather_equals_unsharded():
    L, H, world = 4, 6, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, L, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = torch.tensor([[_fake_ie(l, h) for h in range(H)] for l in range(L)])
    for _, full in results:
        assert torch.equal(full, want)


# ---- the sweep driver under two ranks: rank 0 owns heads_scores.json, the resume state is broadcast ------------------
class _ShardedFakeModel:
    """Stands in for Find_Best_Heads(rank=, world_size=): cal_logits computes this rank's heads and all_gathers."""

    def __init__(self, rank, world, L, H):
        from types import SimpleNamespace
        self.sweep = SimpleNamespace(rank=rank, world_size=world, group=None)
        self.L, self.H, self.calls = L, H, 0

    def cal_logits(self, question="", paras=[], scores=[], wrong_answer=""):
        self.calls += 1
        r, w = self.sweep.rank, self.sweep.world_size
        mine = heads_of_rank(self.H, r, w)
        local = torch.full((self.L, (self.H + w - 1) // w), float("nan"))
        for l in range(self.L):
            for i, h in enumerate(mine):
                local[l, i] = _fake_ie(l, h) + len(question)
        return gather_ie(local, self.L, self.H, w).tolist()


def _driver_worker(rank, world, port, workdir, q):
    import json
    from cram_b200 import find_best_heads as fbh
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    os.chdir(workdir)
    model = _ShardedFakeModel(rank, world, 2, 4)
    out = fbh.casual_tracing_per_head(filepath="data.json", LLM="Meta-Llama-3-8B-Instruct", output_dir="out", model=model)
    dist.barrier()
    q.put((rank, model.calls, out, json.load(open("out/llama3/heads_scores.json"))))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_driver_resumes_in_step_and_only_rank0_writes(tmp_path):
    import json
    L, H, world = 2, 4, 2
    data = [{"question": "q" * (i + 1), "ori_fake": ["fake doc"], "reranked_dense_ctxs": ["a", "b", "c", "d"],
             "wrong answer": "w"} for i in range(3)]
    json.dump(data, open(tmp_path / "data.json", "w"))
    first = [[_fake_ie(l, h) + 1 for h in range(H)] for l in range(L)]
    os.makedirs(tmp_path / "out" / "llama3")
    json.dump([first], open(tmp_path / "out" / "llama3" / "heads_scores.json", "w"))  # sample 0 already done
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_driver_worker, args=(r, world, port, str(tmp_path), q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [[[_fake_ie(l, h) + (i + 1) for h in range(H)] for l in range(L)] for i in range(3)]
    for rank, calls, out, on_disk in results:
        assert calls == 2          # both ranks skipped sample 0 (resume state came from rank 0)
        assert out == want and on_disk == want
    assert not os.path.exists(tmp_path / "out" / "llama3" / "heads_scores.json.tmp")
