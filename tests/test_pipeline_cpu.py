"""Pipeline schedule + neighbour exchange on CPU: world_size 2 and 3 over gloo, with a
toy integer stage standing in for the CUDA ShardEngine (host logic only)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dli_b200.pipeline import PipelineRunner, StageBase, tick_work, total_ticks
from dli_b200.sharding import shard_ranges


class ToyStage(StageBase):
    """hidden' = hidden * 3 + (stage + 1) + step ; last stage: token = hidden.sum % 1000.
    Stage 0 turns tokens into hidden = token (+ prompt sum at prefill)."""

    def __init__(self, rank, world, n_micro, b, d):
        self.rank, self.is_first, self.is_last = rank, rank == 0, rank == world - 1
        self.b, self.d = b, d
        self.tokens = [torch.zeros(b, dtype=torch.int32) for _ in range(n_micro)]
        self.hidden = [torch.zeros(b, d, dtype=torch.float32) for _ in range(n_micro)]
        self.pf_in = torch.zeros(b, d, dtype=torch.float32)
        self.steps = [0] * n_micro
        self.log = [[] for _ in range(n_micro)]
        self.prompts = [torch.arange(b, dtype=torch.float32) + 10 * mb for mb in range(n_micro)]

    def recv_buffer(self, kind, mb):
        if self.is_first:
            return self.tokens[mb]
        return self.pf_in if kind == "prefill" else self.hidden[mb]

    def compute(self, kind, mb):
        if self.is_first:
            src = self.prompts[mb] if kind == "prefill" else self.tokens[mb].float()
            h = src[:, None].expand(self.b, self.d).clone()
        else:
            h = (self.pf_in if kind == "prefill" else self.hidden[mb]).clone()
        h = h * 3 + (self.rank + 1) + self.steps[mb]
        self.steps[mb] += 1
        if self.is_last:
            tok = (h.sum(dim=1).long() % 1000).int()
            self.log[mb].append(tok.clone())
            self.out = tok
            return self.out
        self.hidden[mb].copy_(h)
        return self.hidden[mb]


def sequential(world, n_micro, b, d, rounds):
    """Same toy model without any pipeline: the expected token log per slot."""
    logs = []
    for mb in range(n_micro):
        log, src = [], torch.arange(b, dtype=torch.float32) + 10 * mb
        for step in range(rounds + 1):
            h = src[:, None].expand(b, d).clone()
            for s in range(world):
                h = h * 3 + (s + 1) + step
            tok = (h.sum(dim=1).long() % 1000).int()
            log.append(tok)
            src = tok.float()
        logs.append(torch.stack(log))
    return logs


def _worker(rank, world, port, n_micro, rounds, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stage = ToyStage(rank, world, n_micro, b=2, d=4)
    runner = PipelineRunner(stage, rank, world, n_micro)
    runner.run("prefill")
    runner.run("decode", rounds)
    if rank == world - 1:
        q.put([torch.stack(l).tolist() for l in stage.log])
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,n_micro,rounds", [(2, 2, 3), (2, 1, 2), (3, 4, 2), (2, 3, 4)])
def test_pipeline_matches_sequential_over_gloo(world, n_micro, rounds):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_micro, rounds, q)) for r in range(world)]
    [p.start() for p in procs]
    got = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    want = sequential(world, n_micro, 2, 4, rounds)
    for mb in range(n_micro):
        assert got[mb] == want[mb].tolist(), (mb, got[mb], want[mb].tolist())


def test_tick_schedule_properties():
    for N, M, R in [(2, 2, 3), (4, 4, 5), (8, 8, 2), (3, 5, 2), (4, 2, 3), (2, 1, 4)]:
        P = max(N, M)
        seen = {}
        for t in range(total_ticks(N, M, R)):
            for s in range(N):
                w = tick_work(t, s, N, M, R)
                if w is not None:
                    seen.setdefault(w, []).append((t, s))
        assert len(seen) == M * R  # every (round, slot) is processed ...
        for (r, mb), visits in seen.items():
            assert [s for _, s in visits] == list(range(N))  # ... by every stage in order, one tick apart
            assert [t for t, _ in visits] == list(range(visits[0][0], visits[0][0] + N))
            if r > 0:  # the token of round r-1 has left the last stage before stage 0 needs it
                assert seen[(r - 1, mb)][-1][0] < visits[0][0]
        assert tick_work(total_ticks(N, M, R), 0, N, M, R) is None


def test_stage_layer_ranges_follow_reference_formula():
    assert shard_ranges(32, 4) == [(0, 7), (8, 15), (16, 23), (24, 31)]
    assert shard_ranges(22, 2) == [(0, 10), (11, 21)]
    assert shard_ranges(80, 8)[-1] == (70, 79)
    assert shard_ranges(6, 4) == [(0, 0), (1, 1), (2, 2), (3, 5)]  # remainder on the last shard
    with pytest.raises(ValueError):
        shard_ranges(2, 3)


# ------------------------------------------------------------------ vocabulary-parallel head schedule
class ToyVPStage(ToyStage):
    """ToyStage whose 'head' is split over the stages: stage s proposes
    (value = (sum(hf) * (s + 3)) % 17, id = 100 * s + sum(hf) % 7); the global token is the id of
    the highest value (lowest id on ties)."""

    def __init__(self, rank, world, n_micro, b, d):
        super().__init__(rank, world, n_micro, b, d)
        self.world = world
        self.hf = [torch.zeros(b, d, dtype=torch.float32) for _ in range(n_micro)]
        self.cand = [torch.zeros(b, 2, dtype=torch.float32) for _ in range(n_micro)]
        self.cands = [torch.zeros(world, b, 2, dtype=torch.float32) for _ in range(n_micro)]
        self.final = [[] for _ in range(n_micro)]

    def compute(self, kind, mb):
        if self.is_first:
            src = self.prompts[mb] if kind == "prefill" else self.tokens[mb].float()
            h = src[:, None].expand(self.b, self.d).clone()
        else:
            h = (self.pf_in if kind == "prefill" else self.hidden[mb]).clone()
        h = h * 3 + (self.rank + 1) + self.steps[mb]
        self.steps[mb] += 1
        if self.is_last:
            self.hf[mb].copy_(h)
            return self.hf[mb]
        self.hidden[mb].copy_(h)
        return self.hidden[mb]

    def hf_buffer(self, mb):
        return self.hf[mb]

    @staticmethod
    def propose(hf, s):
        tot = hf.sum(dim=1).long()
        return ((tot * (s + 3)) % 17).float(), (100 * s + tot % 7).float()

    def head_slice(self, mb):
        val, idx = self.propose(self.hf[mb], self.rank)
        tgt = self.cands[mb][0] if self.is_first else self.cand[mb]
        tgt[:, 0], tgt[:, 1] = val, idx
        return tgt

    def cand_buffer(self, src, mb):
        return self.cands[mb][src]

    def combine(self, mb):
        c = self.cands[mb]  # [world, b, 2]
        key = c[:, :, 0] * 10000 - c[:, :, 1]  # highest value, then lowest id
        win = key.argmax(dim=0)
        tok = c[win, torch.arange(self.b), 1].int()
        self.tokens[mb].copy_(tok)
        self.final[mb].append(tok.clone())


def sequential_vp(world, n_micro, b, d, rounds):
    logs = []
    for mb in range(n_micro):
        log, src = [], torch.arange(b, dtype=torch.float32) + 10 * mb
        for step in range(rounds + 1):
            h = src[:, None].expand(b, d).clone()
            for s in range(world):
                h = h * 3 + (s + 1) + step
            vals, ids = zip(*(ToyVPStage.propose(h, s) for s in range(world)))
            key = torch.stack(vals) * 10000 - torch.stack(ids)
            tok = torch.stack(ids)[key.argmax(dim=0), torch.arange(b)].int()
            log.append(tok)
            src = tok.float()
        logs.append(torch.stack(log))
    return logs


def _vp_worker(rank, world, port, n_micro, rounds, q):
    from dli_b200.pipeline import VocabParallelRunner

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stage = ToyVPStage(rank, world, n_micro, b=2, d=4)
    runner = VocabParallelRunner(stage, rank, world, n_micro)
    runner.run("prefill")
    runner.run("decode", rounds)
    if rank == 0:
        q.put([torch.stack(l).tolist() for l in stage.final])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_micro,rounds", [(2, 3, 3), (2, 1, 2), (3, 4, 2), (3, 2, 3)])
def test_vocab_parallel_pipeline_matches_sequential_over_gloo(world, n_micro, rounds):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_vp_worker, args=(r, world, port, n_micro, rounds, q)) for r in range(world)]
    [p.start() for p in procs]
    got = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    want = sequential_vp(world, n_micro, 2, 4, rounds)
    for mb in range(n_micro):
        assert got[mb] == want[mb].tolist(), (mb, got[mb], want[mb].tolist())


# ------------------------------------------------------------------------------ chunked prefill
class ToyChunkStage:
    """Chunk c of slot mb carries the value 100*mb + c; every stage adds 1000*(stage+1) and remembers
    (per slot) the order in which chunks arrived — what a KV cache needs: chunk c after chunks < c."""

    def __init__(self, rank, world, n_micro, d=3):
        self.rank, self.is_first, self.is_last = rank, rank == 0, rank == world - 1
        self.buf = torch.zeros(d)
        self.d = d
        self.seen = [[] for _ in range(n_micro)]

    def chunk_buffer(self):
        return self.buf

    def compute_chunk(self, mb, c):
        h = torch.full((self.d,), float(100 * mb + c)) if self.is_first else self.buf.clone()
        h = h + 1000 * (self.rank + 1)
        self.seen[mb].append((c, h.tolist()))
        self.out = h
        return self.out


def _chunk_worker(rank, world, port, n_micro, n_chunks, q):
    from dli_b200.pipeline import run_chunked_prefill

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stage = ToyChunkStage(rank, world, n_micro)
    run_chunked_prefill(stage, rank, world, n_micro, n_chunks)
    q.put((rank, stage.seen))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_micro,n_chunks", [(2, 3, 4), (3, 2, 3), (2, 1, 1), (3, 4, 2)])
def test_chunked_prefill_schedule_over_gloo(world, n_micro, n_chunks):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_chunk_worker, args=(r, world, port, n_micro, n_chunks, q)) for r in range(world)]
    [p.start() for p in procs]
    got = dict(q.get(timeout=120) for _ in range(world))
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank in range(world):
        for mb in range(n_micro):
            seen = got[rank][mb]
            # chunks 0 .. C-2 of every slot, in order, each carrying what the stages before added
            assert [c for c, _ in seen] == list(range(n_chunks - 1))
            for c, h in seen:
                want = 100 * mb + c + sum(1000 * (s + 1) for s in range(rank + 1))
                assert h == [float(want)] * 3


# ------------------------------------------------------------------------------ worker-group serving loop
def toy_generate(world, first, n):
    """Tokens a ToyMailStage pipeline (b = 1, head on the last stage) consumes, starting from `first`."""
    out, src = [], float(first)
    for step in range(n):
        out.append(int(src))
        h = torch.full((1, 4), src)
        for s in range(world):
            h = h * 3 + (s + 1) + step
        src = float(h.sum(dim=1).long() % 1000)
    return out


def _server_worker(rank, world, port, q):
    """PipelineServer over gloo with a toy data plane: concurrent requests share the pipeline's slots, leave at
    max_length / EOS / deadline, a request no rank can serve fails alone, and the group keeps serving."""
    import threading
    import time

    from dli_b200.pipeline import DistExchange, MailboxRunner, mailbox_period
    from dli_b200.pipeline_worker import PipelineServer, StageBackend

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ctrl = dist.new_group(backend="gloo")
    hop = 2
    P = mailbox_period(1, world, hop, False)

    class Toy(StageBackend):
        n_slots = P

        def __init__(self):
            self.stage = ToyMailStage(rank, world, P, 1, 4, P, hop, False)
            self.stage.consumed = [[] for _ in range(P)]
            orig = self.stage.compute

            def compute(mb):
                if rank == 0:
                    self.stage.consumed[mb].append(int(self.stage.tokens[mb][0]))
                orig(mb)

            self.stage.compute = compute
            self._runner = MailboxRunner(self.stage, rank, world, P, hop=hop, vp=False, exchange=DistExchange())

        def prepare(self, model, probes=None):
            assert probes is not None and len(probes) == world
            if model == "missing" and rank == world - 1:
                raise ValueError(f"Shard {rank} of model missing is not loaded")

        def check_request(self, req):
            if len(req["ids"]) >= req["max_length"]:
                raise ValueError("Input length of input_ids is too long")

        def join(self, slot, req):
            if req["ids"][0] == 666:  # every rank fails at the same point, before any exchange
                raise RuntimeError("KV cache out of pages")
            self.stage.steps[slot] = 0
            self.stage.tokens[slot] = torch.tensor([float(sum(req["ids"]) % 1000)])
            self.stage.consumed[slot] = []

        def runner(self):
            return self._runner

        def read_tokens(self, slot, n):
            if self.stage.consumed[slot][:1] == [777] and not getattr(self, "glitched", False):
                self.glitched = True  # rank 0's bookkeeping fails once, after a block every other rank finished
                raise RuntimeError("bookkeeping")
            return self.stage.consumed[slot][:n]

        def release(self, slot):
            pass

        def reset(self):  # all ranks together: a fresh runner, tick counter back to zero everywhere
            self._runner = MailboxRunner(self.stage, rank, world, P, hop=hop, vp=False, exchange=DistExchange())

    srv = PipelineServer(rank, world, ctrl, Toy(), block=2)
    if rank == 0:
        out = {}

        def client():
            try:
                reqs = {f"r{i}": {"model": "toy", "ids": [i + 1, 7], "max_length": 2 + n, "deadline": None, "gen": {}}
                        for i, n in enumerate([5, 9, 3, 12, 6, 4])}  # six requests on four slots
                eos_tok = toy_generate(world, 101, 20)[6]
                reqs["eos"] = {"model": "toy", "ids": [94, 7], "max_length": 2 + 20, "deadline": None, "gen": {}, "eos": [eos_tok]}
                futs = {k: srv.submit(v) for k, v in reqs.items()}
                late = srv.submit({"model": "toy", "ids": [1, 1], "max_length": 300, "deadline": time.time() - 1, "gen": {}})
                bad = srv.submit({"model": "toy", "ids": [1, 2, 3], "max_length": 3, "deadline": None, "gen": {}})
                for k, f in futs.items():
                    out[k] = (reqs[k], f.result(timeout=100))
                for name, f in (("late", late), ("bad", bad)):
                    try:
                        f.result(timeout=100)
                        out[name] = "no error"
                    except Exception as e:  # noqa: BLE001
                        out[name] = f"{type(e).__name__}: {e}"
                try:
                    srv.submit({"model": "missing", "ids": [5], "max_length": 9, "deadline": None, "gen": {}}).result(timeout=100)
                except Exception as e:  # noqa: BLE001
                    out["missing"] = str(e)
                r = {"model": "toy", "ids": [40, 2], "max_length": 6, "deadline": None, "gen": {}}
                out["after"] = (r, srv.submit(r).result(timeout=100))  # the group still serves
                # a block that fails (on every rank in a join; on rank 0 alone after the ticks): what was in flight fails,
                # every rank rebuilds its stage with the next plan, and later requests get the right tokens again
                for name, ids in (("boom", [666, 0]), ("glitch", [777, 0])):
                    side = srv.submit({"model": "toy", "ids": [3, 3], "max_length": 40, "deadline": None, "gen": {}})
                    try:
                        srv.submit({"model": "toy", "ids": ids, "max_length": 12, "deadline": None, "gen": {}}).result(timeout=100)
                        out[name] = "no error"
                    except Exception as e:  # noqa: BLE001
                        out[name] = f"{type(e).__name__}: {e}"
                    try:
                        side.result(timeout=100)
                        out[name + "_side"] = "no error"
                    except Exception as e:  # noqa: BLE001
                        out[name + "_side"] = f"{type(e).__name__}: {e}"
                    r = {"model": "toy", "ids": [41 + len(name), 2], "max_length": 11, "deadline": None, "gen": {}}
                    out["after_" + name] = (r, srv.submit(r).result(timeout=100))
                out["stats"] = dict(srv.stats)
            finally:
                srv.shutdown()

        th = threading.Thread(target=client)
        th.start()
        srv.serve_forever()
        th.join()
        q.put(out)
    else:
        srv.serve_forever()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_worker_group_serves_concurrent_requests_over_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_server_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    out = q.get(timeout=180)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for k, v in out.items():
        if k in ("late", "bad", "missing", "stats", "boom", "glitch", "boom_side", "glitch_side"):
            continue
        req, toks = v
        n = req["max_length"] - len(req["ids"])
        want = toy_generate(world, sum(req["ids"]) % 1000, n)
        if req.get("eos"):
            want = want[: want.index(req["eos"][0]) + 1]
            assert len(want) == 7
        assert toks == want, (k, toks, want)
    assert out["late"] == "TimeoutError: Inference generation timed out"
    assert out["bad"].startswith("ValueError: Input length of input_ids")
    assert out["missing"] == f"Shard {world - 1} of model missing is not loaded"
    assert out["boom"] == "RuntimeError: KV cache out of pages" and out["glitch"] == "RuntimeError: bookkeeping"
    # the request that shared the failed block fails with it (its slot's state is gone), never with wrong tokens
    assert out["boom_side"].startswith("RuntimeError") and out["glitch_side"].startswith("RuntimeError")
    assert out["stats"]["max_active"] >= 3  # requests really shared the pipeline


# ------------------------------------------------------------------------------ block-wise decode (return_last)
def _block_worker(rank, world, port, n_micro, blocks, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stage = ToyStage(rank, world, n_micro, b=2, d=4)
    runner = PipelineRunner(stage, rank, world, n_micro, return_last=True)
    runner.run("prefill")
    for k in blocks:  # a server drives the generation as several run("decode", k) calls
        runner.run("decode", k)
    if rank == world - 1:
        q.put([torch.stack(l).tolist() for l in stage.log])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_micro,blocks", [(2, 1, [2, 1, 3]), (3, 1, [1, 1, 4]), (2, 2, [3, 2])])
def test_decode_in_blocks_equals_one_run(world, n_micro, blocks):
    """return_last=True: the last round's tokens of every run() travel back to stage 0, so consecutive
    run("decode", k) calls continue the same generation (what pipeline_worker.py relies on)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_block_worker, args=(r, world, port, n_micro, blocks, q)) for r in range(world)]
    [p.start() for p in procs]
    got = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    want = sequential(world, n_micro, 2, 4, sum(blocks))
    for mb in range(n_micro):
        assert got[mb] == want[mb].tolist(), (mb, got[mb], want[mb].tolist())


def test_chunk_plan_covers_the_prompt_exactly():
    from dli_b200.pipeline import chunk_plan

    for T in list(range(1, 70)) + [2048, 8192, 8191, 4097]:
        for want in (1, 2, 3, 4, 8, 100):
            chunks, tc, last = chunk_plan(T, want)
            assert 1 <= chunks <= max(1, min(want, T)) and 1 <= last <= tc
            assert (chunks - 1) * tc + last == T
    assert chunk_plan(8192, 8) == (8, 1024, 1024)
    assert chunk_plan(10, 3) == (3, 4, 2)
    assert chunk_plan(5, 1) == (1, 5, 5)


# ------------------------------------------------------------------------------ mailbox pipeline (one all-gather per tick)
class ToyMailStage:
    """The toy model of ToyStage / ToyVPStage in mailbox form: mail = [hidden b*d | candidate b*2 | token b] floats."""

    def __init__(self, rank, world, n_micro, b, d, P, H, vp):
        self.rank, self.N, self.b, self.d, self.P, self.H, self.vp = rank, world, b, d, P, H, vp
        self.is_first, self.is_last = rank == 0, rank == world - 1
        L = b * d + 3 * b
        self.out = torch.zeros(P, L)
        self.inn = torch.zeros(P, world, L)
        self.tokens = [torch.arange(b, dtype=torch.float32) + 10 * mb for mb in range(n_micro)]  # first inputs
        self.steps = [0] * n_micro
        self.final = [[] for _ in range(n_micro)]   # stage 0: tokens taken in
        self.made = [[] for _ in range(n_micro)]    # last stage without vp: tokens produced

    def mail_out(self, ph):
        return self.out[ph]

    def mail_in(self, ph):
        return self.inn[ph]

    def _h(self, mail):
        return mail[: self.b * self.d].view(self.b, self.d)

    def compute(self, mb):
        ph = (mb + self.H * self.rank) % self.P
        rd = (ph - self.H) % self.P
        h = (self.tokens[mb][:, None].expand(self.b, self.d) if self.is_first else self._h(self.inn[rd][self.rank - 1])).clone()
        h = h * 3 + (self.rank + 1) + self.steps[mb]
        self.steps[mb] += 1
        if self.is_last and not self.vp:
            tok = (h.sum(dim=1).long() % 1000).float()
            self.out[ph][self.b * self.d + 2 * self.b:] = tok
            self.made[mb].append(tok.int().clone())
        else:
            self._h(self.out[ph]).copy_(h)

    def head(self, mb):
        ph = (mb + self.H * self.N) % self.P
        hf = self._h(self.inn[(ph - self.H) % self.P][self.N - 1])
        val, idx = ToyVPStage.propose(hf, self.rank)
        c = self.out[ph][self.b * self.d: self.b * self.d + 2 * self.b].view(self.b, 2)
        c[:, 0], c[:, 1] = val, idx

    def ingest(self, mb):
        sig = self.N + 1 if self.vp else self.N
        rd = ((mb + self.H * sig) % self.P - self.H) % self.P
        if self.vp:
            c = self.inn[rd][:, self.b * self.d: self.b * self.d + 2 * self.b].reshape(self.N, self.b, 2)
            key = c[:, :, 0] * 10000 - c[:, :, 1]
            tok = c[key.argmax(dim=0), torch.arange(self.b), 1].clone()
        else:
            tok = self.inn[rd][self.N - 1][self.b * self.d + 2 * self.b:].clone()
        self.tokens[mb] = tok
        self.final[mb].append(tok.int().clone())


def sequential_mail(world, mb, b, d, rounds, vp):
    log, src = [], torch.arange(b, dtype=torch.float32) + 10 * mb
    for step in range(rounds):
        h = src[:, None].expand(b, d).clone()
        for s in range(world):
            h = h * 3 + (s + 1) + step
        if vp:
            vals, ids = zip(*(ToyVPStage.propose(h, s) for s in range(world)))
            key = torch.stack(vals) * 10000 - torch.stack(ids)
            tok = torch.stack(ids)[key.argmax(dim=0), torch.arange(b)]
        else:
            tok = (h.sum(dim=1).long() % 1000).float()
        log.append(tok.int().tolist())
        src = tok.float()
    return log


def _mail_worker(rank, world, port, n_micro, rounds, hop, vp, dynamic, q):
    from dli_b200.pipeline import DistExchange, MailboxRunner, mailbox_period

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = mailbox_period(n_micro, world, hop, vp)
    stage = ToyMailStage(rank, world, n_micro, 2, 4, P, hop, vp)
    runner = MailboxRunner(stage, rank, world, n_micro, hop=hop, vp=vp, exchange=DistExchange())
    assert runner.P == P
    if not dynamic:
        runner.run(rounds)
        runner.run(1)  # a second run continues every slot's generation from the drained pipeline
        want_rounds = [rounds + 1] * n_micro
    else:
        # slots join at different ticks and one leaves early, as requests do in the serving loop
        runner.open_slot(0)
        runner.ticks(3)
        for j in range(1, n_micro):
            runner.open_slot(j)
        runner.ticks(2 * P + 1)
        runner.close_slot(0)
        runner.ticks(2 * P)
        for j in range(1, n_micro):
            runner.close_slot(j)
        while runner.busy():
            runner.tick()
        runner.flush()
        # rounds every slot was open for: its stage-0 ticks inside [begin, end)
        want_rounds = [(runner.end[j] - runner.begin[j] + P - 1) // P for j in range(n_micro)]
    if rank == 0:
        q.put(([torch.stack(l).tolist() if l else [] for l in stage.final], want_rounds))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_micro,rounds,hop,vp,dynamic", [
    (2, 3, 3, 1, True, False), (2, 6, 3, 2, True, False), (3, 2, 2, 2, True, False), (3, 8, 2, 2, True, False),
    (2, 2, 3, 1, False, False), (3, 6, 2, 2, False, False), (2, 1, 4, 2, False, False),
    (2, 6, 0, 2, True, True), (3, 4, 0, 2, False, True), (2, 3, 0, 1, True, True),
])
def test_mailbox_pipeline_matches_sequential_over_gloo(world, n_micro, rounds, hop, vp, dynamic):
    """One all-gather of fixed-size mails per tick replaces every pairwise send: tokens equal the sequential toy model
    for hop 1 (serialised) and hop 2 (exchange under the next tick), with the head split over the stages (vp) or on the
    last stage, for fixed runs and for slots that join and leave between ticks."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_mail_worker, args=(r, world, port, n_micro, rounds, hop, vp, dynamic, q))
             for r in range(world)]
    [p.start() for p in procs]
    got, want_rounds = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for mb in range(n_micro):
        n = want_rounds[mb] if want_rounds else len(got[mb])
        assert n > 0 and len(got[mb]) == n
        assert got[mb] == sequential_mail(world, mb, 2, 4, n, vp), (mb, got[mb])
    if dynamic:
        assert len(set(want_rounds)) > 1 or n_micro <= 3  # the scenario really has slots of different lifetimes


def test_mailbox_schedule_every_piece_reads_what_was_delivered():
    """Pure schedule check: with P = max(M, H*(N+1)) the mail a piece reads at tick t (phase (t-H) mod P) is never
    overwritten between its delivery and its use, for every stage offset."""
    from dli_b200.pipeline import mailbox_period

    for N in (1, 2, 4, 8):
        for H in (1, 2):
            for vp in (True, False):
                for M in (1, N, 2 * N + 2, 20):
                    P = mailbox_period(M, N, H, vp)
                    assert P >= M and P >= H * (N + 1 if vp else N) and P > H or N == 1
                    # tick t writes phase t % P; the next write of that phase is tick t + P > t + H (its reader)
                    assert P > H or (N == 1 and not vp and M == 1)
