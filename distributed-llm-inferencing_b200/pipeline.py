"""Layer-range pipeline over torch.distributed point-to-point (NCCL over NVLink on
the B200 box, gloo in the CPU tests).

This is the reference's DECLARED design — a model split into N contiguous layer
ranges on N nodes (shard_model.py:55-66) whose shards pass hidden states along
(worker/app.py:332-335 comment, never implemented: SURVEY.md §0 F1) — with the
HTTP/JSON hop replaced by send/recv of the `[b, d]` bf16 hidden tile between
neighbouring GPUs, and the sampled `[b]` int32 tokens returned from the last
stage to stage 0.  There is no collective on this path (no tensor parallelism):
only neighbour exchanges, issued as ONE grouped batch per tick so the forward
sends, the token return and the matching receives progress together.

Schedule.  Micro-batch slots mb = 0..M-1 (each `b` sequences) are in flight.
With period P = max(M, N), stage s works at tick t on u = t - s:
    round r = u // P,  slot mb = u % P   (idle if mb >= M or r outside the run)
so the token of (r, mb) leaves the last stage at tick r*P + mb + N - 1 and reaches
stage 0 before tick (r+1)*P + mb.  Every tick ends with the grouped exchange:
send what was just computed, receive the input of the next tick.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist


def tick_work(t: int, stage: int, n_stages: int, n_micro: int, n_rounds: int):
    """(round, slot) stage `stage` works on at tick `t`, or None when idle."""
    P = max(n_micro, n_stages)
    u = t - stage
    if u < 0:
        return None
    r, mb = divmod(u, P)
    if mb >= n_micro or r >= n_rounds:
        return None
    return r, mb


def total_ticks(n_stages: int, n_micro: int, n_rounds: int) -> int:
    P = max(n_micro, n_stages)
    return (n_rounds - 1) * P + n_micro + n_stages - 1


class StageBase:
    """What the schedule needs from a stage.  `recv_buffer(kind, mb)` is where the
    stage's input for slot mb lands (hidden state; on stage 0 the returned tokens);
    `compute(kind, mb)` consumes it and returns the tensor to pass on (hidden state;
    on the last stage the tokens)."""

    is_first: bool
    is_last: bool

    def recv_buffer(self, kind: str, mb: int) -> torch.Tensor:
        raise NotImplementedError

    def compute(self, kind: str, mb: int) -> torch.Tensor:
        raise NotImplementedError


class PipelineRunner:
    def __init__(self, stage: StageBase, rank: int, world: int, n_micro: int, group=None, return_last: bool = False):
        """return_last: also send the tokens of the LAST decode round of a run() back to stage 0, so that a
        generation can be driven as several run("decode", k) calls (a server checking EOS / its deadline
        between blocks); a single run() does not need them back."""
        self.stage, self.rank, self.world, self.M, self.group = stage, rank, world, n_micro, group
        self.return_last = return_last

    def _exchange(self, t: int, kind: str, n_rounds: int, mine, out: Optional[torch.Tensor], token_return: bool):
        """Grouped send of this tick's result and receive of the next tick's input."""
        N, s = self.world, self.rank
        ops: List[dist.P2POp] = []
        if mine is not None:
            r, mb = mine
            if not self.stage.is_last:
                ops.append(dist.P2POp(dist.isend, out, s + 1, self.group))
            elif token_return and N > 1 and (kind == "prefill" or r + 1 < n_rounds or self.return_last):
                ops.append(dist.P2POp(dist.isend, out, 0, self.group))
        if s > 0:
            prev = tick_work(t, s - 1, N, self.M, n_rounds)
            if prev is not None:
                ops.append(dist.P2POp(dist.irecv, self.stage.recv_buffer(kind, prev[1]), s - 1, self.group))
        elif token_return and N > 1:
            last = tick_work(t, N - 1, N, self.M, n_rounds)
            if last is not None and (kind == "prefill" or last[0] + 1 < n_rounds or self.return_last):
                ops.append(dist.P2POp(dist.irecv, self.stage.recv_buffer("decode", last[1]), N - 1, self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()

    def run(self, kind: str, n_rounds: int = 1, token_return: bool = True) -> None:
        """kind='prefill': one pass of every slot's prompt through the stages (the first
        tokens end up in stage 0's decode input buffers).  kind='decode': n_rounds tokens
        for every slot."""
        if kind == "prefill":
            n_rounds = 1
        for t in range(total_ticks(self.world, self.M, n_rounds)):
            mine = tick_work(t, self.rank, self.world, self.M, n_rounds)
            out = self.stage.compute(kind, mine[1]) if mine is not None else None
            self._exchange(t, kind, n_rounds, mine, out, token_return)


def run_chunked_prefill(stage, rank: int, world: int, n_micro: int, n_chunks: int, group=None) -> None:
    """Chunked prefill ("chunk the prefill to pipeline it").  A prompt pass through N
    stages keeps only one stage busy at a time; cut into C chunks of tokens, stage s can work on chunk
    c+1 while stage s+1 works on chunk c (attention reads the earlier chunks' keys/values from the paged
    cache).  This runs chunks 0 .. C-2 of every slot — unit u = slot*(C-1) + chunk, stage s works on unit
    t - s at tick t, one grouped send/receive per tick; the LAST chunk of every slot then goes through
    the runner's ordinary "prefill" pass, which ends with the head and the first token.
    stage.compute_chunk(slot, chunk) -> hidden states to pass on; stage.chunk_buffer() -> where the
    next unit's input lands."""
    units = n_micro * (n_chunks - 1)
    if units <= 0:
        return
    for t in range(units + world - 1):
        u = t - rank
        out = None
        if 0 <= u < units:
            mb, c = divmod(u, n_chunks - 1)
            out = stage.compute_chunk(mb, c)
        ops: List[dist.P2POp] = []
        if out is not None and rank < world - 1:
            ops.append(dist.P2POp(dist.isend, out, rank + 1, group))
        if rank > 0 and 0 <= t - (rank - 1) < units:
            ops.append(dist.P2POp(dist.irecv, stage.chunk_buffer(), rank - 1, group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()


# ------------------------------------------------------------------------------ vocabulary-parallel head
class VocabParallelRunner:
    """Same pipeline, but the lm_head is split by vocabulary rows over ALL stages.

    With the whole head on the last stage that stage streams its layers plus the [V, d]
    matrix every tick (Llama-3-8B on 8 GPUs: 1.74 GB + 1.05 GB against 1.74 GB elsewhere) and
    sets the tick for everybody.  Here every stage holds V/N rows (the reference's sharder
    copies the head into every shard anyway) and a slot takes two more ticks per round:

      tick T+s      stage s runs its layers on the slot              (s = 0..N-1)
      end of T+N-1  the last stage sends the slot's final hidden state to every stage
      tick T+N      every stage: final norm + its head slice -> (best logit, token id)
      end of T+N    stages 1..N-1 send their candidate to stage 0
      tick T+N+1    stage 0 takes the global argmax (lowest id on ties) = next input token

    so the period is P = max(M, N+1) and all stages stay busy with M >= N+1 slots.  Every
    tick still ends with ONE grouped batch of sends/receives (per pair of ranks the sends and
    the matching receives are posted in the same order: hidden state, final hidden, candidate).
    Greedy only: the token is the same one a single argmax over the full row would pick."""

    def __init__(self, stage, rank: int, world: int, n_micro: int, group=None):
        self.stage, self.rank, self.world, self.M, self.group = stage, rank, world, n_micro, group
        self.P = max(n_micro, world + 1)

    def _idx(self, u: int, n_rounds: int):
        if u < 0:
            return None
        r, mb = divmod(u, self.P)
        return (r, mb) if (mb < self.M and r < n_rounds) else None

    def total_ticks(self, n_rounds: int) -> int:
        return (n_rounds - 1) * self.P + self.M + self.world + 1

    def run(self, kind: str, n_rounds: int = 1, on_tokens=None) -> None:
        """on_tokens(round, slot, tokens) is called on stage 0 right after a slot's next tokens
        are known (e.g. to stream them to the host)."""
        if kind == "prefill":
            n_rounds = 1
        N, s, st = self.world, self.rank, self.stage
        for t in range(self.total_ticks(n_rounds)):
            if s == 0:
                jc = self._idx(t - N - 1, n_rounds)
                if jc is not None:
                    toks = st.combine(jc[1])
                    if on_tokens is not None:
                        on_tokens(jc[0], jc[1], toks)
            j = self._idx(t - s, n_rounds)
            out = st.compute(kind, j[1]) if j is not None else None
            jh = self._idx(t - N, n_rounds)
            cand = st.head_slice(jh[1]) if jh is not None else None
            ops: List[dist.P2POp] = []
            if j is not None and s < N - 1:
                ops.append(dist.P2POp(dist.isend, out, s + 1, self.group))
            if s > 0:
                jp = self._idx(t - (s - 1), n_rounds)
                if jp is not None:
                    ops.append(dist.P2POp(dist.irecv, st.recv_buffer(kind, jp[1]), s - 1, self.group))
            jl = self._idx(t - (N - 1), n_rounds)
            if jl is not None and N > 1:
                if s == N - 1:
                    for peer in range(N - 1):
                        ops.append(dist.P2POp(dist.isend, out, peer, self.group))
                else:
                    ops.append(dist.P2POp(dist.irecv, st.hf_buffer(jl[1]), N - 1, self.group))
            if jh is not None and N > 1:
                if s > 0:
                    ops.append(dist.P2POp(dist.isend, cand, 0, self.group))
                else:
                    for peer in range(1, N):
                        ops.append(dist.P2POp(dist.irecv, st.cand_buffer(peer, jh[1]), peer, self.group))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()


# ------------------------------------------------------------------------------ mailbox pipeline (decode)
class DistExchange:
    """The per-tick exchange over torch.distributed: ONE all-gather of every stage's outgoing mail
    (NCCL over NVLink on the B200 box, gloo in the CPU tests).  With a `comm_stream` the collective is issued on
    that stream behind an event of the compute stream, so it runs while the compute stream goes on with the next
    tick; wait() makes the current stream wait for it."""

    def __init__(self, group=None, comm_stream=None):
        self.group, self.comm = group, comm_stream

    def issue(self, mail_out: torch.Tensor, mail_in: torch.Tensor, phase=None):
        if self.comm is None:
            return dist.all_gather_into_tensor(mail_in.view(-1), mail_out, group=self.group, async_op=True)
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(ev)
            return dist.all_gather_into_tensor(mail_in.view(-1), mail_out, group=self.group, async_op=True)

    def wait(self, work) -> None:
        work.wait()


class LocalExchange:
    """The same exchange between N stages living in ONE process (tests: N layer ranges on one GPU, driven tick by
    tick in rank order): issue() copies the mail straight into every stage's incoming mailbox."""

    def __init__(self):
        self.stages = []  # filled by the caller: stage objects in rank order

    def bind(self, rank: int):
        ex = self

        class _Port:
            def issue(self, mail_out, mail_in, phase=None):
                for st in ex.stages:
                    st.mail_in(phase)[rank].copy_(mail_out)
                return None

            def wait(self, work):
                pass

        return _Port()


class MailboxRunner:
    """Decode pipeline over per-tick mailboxes.

    Stage s works on slot u mod P at tick t = u + H*s (H = `hop` ticks between neighbouring stages); the head
    slices of a vocabulary-parallel head (vp) run at stage offset N and stage 0 takes the slot's next token in at
    offset N+1 (vp) or N (whole head on the last stage), so a slot's round lasts P = max(M, H*(N+1)) or
    max(M, H*N) ticks.  Every tick ends with one exchange (all-gather of the stages' mails, see
    ShardEngine.attach_mailbox); tick t reads what the exchange of tick t-H delivered.  H = 1 is the serialised
    schedule (the next tick waits for the exchange it has just issued); H = 2 leaves a whole tick for the exchange,
    which then runs on its own stream UNDER the next tick's kernels — the hand-off disappears from the critical
    path at the price of twice the slots in flight.

    `stage` needs: is_first; mail_out(phase), mail_in(phase) -> tensors; compute(slot); head(slot); ingest(slot).
    Slots become active with open()/run() and stay so until close_slot()/the run's last round: the serving loop
    (pipeline_worker.py) lets requests join and leave between ticks, the benchmark times whole periods of a full
    pipeline."""

    def __init__(self, stage, rank: int, world: int, n_micro: int, hop: int = 2, vp: bool = True, exchange=None):
        self.stage, self.rank, self.N, self.M, self.H, self.vp = stage, rank, world, n_micro, hop, vp
        self.sig_head = world
        self.sig_in = world + 1 if vp else world
        self.P = max(n_micro, hop * self.sig_in)
        self.ex = exchange if exchange is not None else DistExchange()
        self.t = 0
        self.pending = {}
        # slot j is active for stage-0 ticks u = j (mod P) with begin[j] <= u < end[j]
        self.begin = [None] * n_micro
        self.end = [None] * n_micro

    # -- slot windows -------------------------------------------------------------------------------------------
    def next_u(self, slot: int, not_before: int = None) -> int:
        """First stage-0 tick >= not_before (default: now) that belongs to `slot`."""
        t0 = self.t if not_before is None else not_before
        return t0 + (slot - t0) % self.P

    def open_slot(self, slot: int, rounds: int = None) -> int:
        """Slot's first round starts at its next stage-0 tick; `rounds` None = until close_slot().  Every rank
        must make the same calls at the same tick.  Returns the tick of the slot's first stage-0 step."""
        u0 = self.next_u(slot)
        self.begin[slot] = u0
        self.end[slot] = None if rounds is None else u0 + rounds * self.P
        return u0

    def close_slot(self, slot: int) -> int:
        """No round of `slot` starts at or after now; rounds already under way finish.  Returns the tick by which
        the slot's last token has been taken in by stage 0."""
        if self.begin[slot] is None:
            return self.t
        self.end[slot] = max(self.begin[slot], self.next_u(slot))
        return self.end[slot] - self.P + self.H * self.sig_in + 1

    def _active(self, sigma: int, t: int):
        u = t - self.H * sigma
        if u < 0:
            return None
        j = u % self.P
        if j >= self.M or self.begin[j] is None or u < self.begin[j] or (self.end[j] is not None and u >= self.end[j]):
            return None
        return j

    def busy(self) -> bool:
        """Is any piece of any open round still to run at or after the current tick?"""
        for j in range(self.M):
            if self.begin[j] is None:
                continue
            if self.end[j] is None:
                return True
            last_u = self.end[j] - self.P  # stage-0 tick of the slot's last round (< begin: no round at all)
            if last_u >= self.begin[j] and last_u + self.H * self.sig_in >= self.t:
                return True
        return False

    # -- one tick ------------------------------------------------------------------------------------------------
    def tick(self) -> None:
        t, st = self.t, self.stage
        w = self.pending.pop(t - self.H, None)
        if w is not None:
            self.ex.wait(w)  # the exchange issued H ticks ago carries this tick's inputs
        ji = self._active(self.sig_in, t) if st.is_first else None
        jc = self._active(self.rank, t)
        jh = self._active(self.sig_head, t) if self.vp else None
        if hasattr(st, "tick_all"):
            st.tick_all(ji, jc, jh)  # one submission for the whole tick
        else:
            if ji is not None:
                st.ingest(ji)
            if jc is not None:
                st.compute(jc)
            if jh is not None:
                st.head(jh)
        ph = t % self.P
        self.pending[t] = self.ex.issue(st.mail_out(ph), st.mail_in(ph), phase=ph)
        self.t += 1

    def ticks(self, n: int) -> None:
        for _ in range(n):
            self.tick()

    def flush(self) -> None:
        """Wait for every exchange still in flight (before the host reads results, or before a barrier)."""
        for t in sorted(self.pending):
            w = self.pending.pop(t)
            if w is not None:
                self.ex.wait(w)

    def run(self, n_rounds: int) -> None:
        """n_rounds tokens for every slot, from a drained pipeline to a drained pipeline (fill and drain included)."""
        for j in range(self.M):
            self.open_slot(j, n_rounds)
        while self.busy():
            self.tick()
        self.flush()


class MailStage:
    """A ShardEngine in mailbox mode as the runner's stage."""

    def __init__(self, engine, b: int, n_micro: int, runner_period: int, hop: int, rank: int, world: int,
                 with_logits: bool = False):
        self.eng, self.b = engine, b
        self.is_first = engine.is_first
        self.sampling = set()  # slots whose request samples (stage 0 then draws from the whole logits row)
        engine.attach_mailbox(b, n_micro, runner_period, hop, rank, world, with_logits=with_logits)
        for mb in range(n_micro):
            engine._make_state(b, mb)

    def mail_out(self, phase):
        return self.eng._mail["out"][phase]

    def mail_in(self, phase):
        return self.eng._mail["in"][phase]

    def compute(self, slot):
        self.eng.mail_compute(slot)

    def head(self, slot):
        self.eng.mail_head(slot)

    def ingest(self, slot):
        self.eng.mail_ingest(slot)

    def tick_all(self, ingest, compute, head):
        self.eng.mail_tick(ingest, compute, head, sample=ingest is not None and ingest in self.sampling)


def mailbox_period(n_micro: int, world: int, hop: int, vp: bool) -> int:
    return max(n_micro, hop * (world + 1 if vp else world))


def chunk_plan(prompt_len: int, want_chunks: int):
    """(chunks, tokens per chunk, tokens in the last chunk) for a prompt fed in about `want_chunks` chunks:
    equal chunks of ceil(T / want) tokens, the last one takes the rest (never empty)."""
    chunks = max(1, min(int(want_chunks), int(prompt_len)))
    tc = (prompt_len + chunks - 1) // chunks
    chunks = (prompt_len + tc - 1) // tc
    return chunks, tc, prompt_len - (chunks - 1) * tc


# ------------------------------------------------------------------------------ CUDA stage
class EngineStage(StageBase):
    """A ShardEngine as one pipeline stage: slot mb <-> the engine's micro-batch state."""

    def __init__(self, engine, b: int, prompt_len: int, max_total_len: int, n_micro: int, prefill_chunks: int = 1):
        self.eng, self.b, self.T = engine, b, prompt_len
        self.is_first, self.is_last = engine.is_first, engine.is_last
        self.prompt_ids = {}  # stage 0: slot -> int32 [b*T] device tensor
        for mb in range(n_micro):
            engine.start_batch(b, prompt_len, max_total_len, mb=mb)
        dev = engine.device
        # chunked prefill: chunks 0..C-2 hold Tc tokens of every prompt, the last one the rest
        self.chunks, self.Tc, self.T_last = chunk_plan(prompt_len, prefill_chunks)
        self._pf_in = None if self.is_first else torch.empty(b * self.Tc, engine.spec.d_model,
                                                             dtype=torch.bfloat16, device=dev)

    def recv_buffer(self, kind, mb):
        st = self.eng._state(self.b, mb)
        if self.is_first:
            return st["tokens"]
        return self._pf_in[:self.b * self.T_last] if kind == "prefill" else st["h"]

    def _chunk_ids(self, mb, c0, n):
        return self.prompt_ids[mb].view(self.b, self.T)[:, c0:c0 + n].reshape(-1).contiguous()

    def chunk_buffer(self):
        return self._pf_in

    def compute_chunk(self, mb, c):
        """Chunk c < chunks-1 of slot mb (run_chunked_prefill); returns the hidden states to pass on."""
        kw = {"ids": self._chunk_ids(mb, c * self.Tc, self.Tc)} if self.is_first else {"hidden": self._pf_in}
        return self.eng.prefill(self.b, self.Tc, mb=mb, prefix=c * self.Tc, final=False, **kw)

    def compute(self, kind, mb):
        if kind == "prefill":
            if self.chunks > 1:  # the last chunk: ends with the head like an unchunked prefill
                p0 = (self.chunks - 1) * self.Tc
                kw = ({"ids": self._chunk_ids(mb, p0, self.T_last)} if self.is_first
                      else {"hidden": self._pf_in[:self.b * self.T_last]})
                return self.eng.prefill(self.b, self.T_last, mb=mb, prefix=p0, **kw)
            if self.is_first:
                return self.eng.prefill(self.b, self.T, ids=self.prompt_ids[mb], mb=mb)
            return self.eng.prefill(self.b, self.T, hidden=self._pf_in, mb=mb)
        out = self.eng.decode_step(self.b, mb=mb)
        if self.is_last and self.eng.head_shard is not None:
            hf = self.eng._state(self.b, mb)["hf"]  # the slot's final hidden, read by this stage's head slice
            hf.copy_(out)
            return hf
        return out

    # vocabulary-parallel head (VocabParallelRunner)
    def hf_buffer(self, mb):
        return self.eng._state(self.b, mb)["hf"]

    def head_slice(self, mb):
        return self.eng.head_slice_step(self.b, mb)

    def cand_buffer(self, src, mb):
        return self.eng._state(self.b, mb)["cands"][src]

    def combine(self, mb):
        return self.eng.combine_candidates(self.b, mb)


def init_distributed():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    return rank, world, local


def build_stage_engine(spec, rank: int, world: int, local: int, max_batch: int, max_seq_len: int, page_size: int = 16,
                       vocab_parallel: bool = False):
    from .engine import ShardEngine
    from .sharding import shard_ranges

    a, b = shard_ranges(spec.n_layers, world)[rank]
    return ShardEngine(spec, a, b, device=f"cuda:{local}", max_batch=max_batch, max_seq_len=max_seq_len,
                       page_size=page_size, head_shard=(rank, world) if vocab_parallel else None)


def _timed(fn):
    """Device time of fn() in ms, max over ranks, bracketed by barriers and synchronisation on both sides."""
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    return t.item()


def measure_pipeline(spec, model_name, rank, world, local, b, n_slots, ctx, K, W, page_size, peaks, vp=True, hop=2,
                     prompt_seed=1, parity_tokens=0, timed_prefill=True):
    """One pipeline configuration: `world` stages (reference sharder ranges), n_slots micro-batch slots of b
    sequences, every sequence with `ctx` tokens of context in the middle of the timed region.  Prefill over
    send/recv (chunked for long prompts on many stages), decode over per-tick mailboxes (MailboxRunner, `hop`).
    K whole periods of the FULL pipeline are timed (no fill or drain inside the timed region).
    parity_tokens > 0: rank 0 also runs the single full-range engine on every slot's prompt, teacher-forced along
    the pipeline's first tokens, and raises unless the pipeline picked the single engine's token at every step
    (or one whose logit ties with it to the last bf16 bit)."""
    from . import ops
    from .measure import ClockSampler, decode_step_bytes

    M, N = n_slots, world
    P = mailbox_period(M, N, hop, vp)
    warm_rounds = max(W, 3, parity_tokens)
    prompt_len = max(8, ctx - warm_rounds - K // 2)  # so that the context is `ctx` in the middle of the timed region
    max_len = prompt_len + warm_rounds + 2 * K + 16  # warm-up, the timed rounds, the host-streamed rounds, drain
    eng = build_stage_engine(spec, rank, world, local, M * b, max_len, page_size, vocab_parallel=vp)
    eng.load_random(seed=0)
    # chunked prefill: stages work on different chunks of a prompt at once.  It pays for long prompts on many stages
    # (70B, 8 stages, 8k tokens, 8 chunks: 8.8 -> 5.5 s for 36 prompts), not at 2k tokens on 2 stages.
    chunks = int(os.environ.get("DLI_PREFILL_CHUNKS", "8" if (world >= 4 and prompt_len >= 4000) else "1"))
    stage = EngineStage(eng, b, prompt_len, max_len, M, prefill_chunks=chunks)
    g = torch.Generator().manual_seed(prompt_seed)
    prompts = [torch.randint(0, spec.vocab, (b * prompt_len,), generator=g, dtype=torch.int32) for _ in range(M)]
    if stage.is_first:
        for mb in range(M):
            stage.prompt_ids[mb] = prompts[mb].cuda()
    pf_runner = (VocabParallelRunner if vp else PipelineRunner)(stage, rank, world, M)

    def prefill_all():
        run_chunked_prefill(stage, rank, world, M, stage.chunks)  # chunks 0..C-2 of every slot (nothing if C == 1)
        pf_runner.run("prefill")  # last chunk + head + first token (lands in stage 0's per-slot token buffer)

    if timed_prefill:
        prefill_all()  # warm-up: lazy module loads, function attributes, NCCL connections
        for mb in range(M):
            eng.start_batch(b, prompt_len, max_len, mb=mb)
        pf_ms = _timed(prefill_all)
    else:
        pf_ms = _timed(prefill_all)  # a single pass (first-call costs included); used for the large named shapes
    firsts = None
    if not vp and eng.is_last:
        firsts = torch.stack([eng._state(b, mb)["tokens"].clone() for mb in range(M)])  # [M, b]

    ms_stage = MailStage(eng, b, M, P, hop, rank, world)
    comm = torch.cuda.Stream() if hop > 1 else None
    runner = MailboxRunner(ms_stage, rank, world, M, hop=hop, vp=vp, exchange=DistExchange(comm_stream=comm))
    for j in range(M):
        runner.open_slot(j)
    fill = hop * (N + 1 if vp else N)
    runner.ticks(warm_rounds * P + fill)  # fills the pipeline, captures every slot's graphs, warm-up rounds
    runner.flush()

    parity = None
    if parity_tokens:
        parity = _parity_against_single_engine(spec, eng, rank, world, local, b, M, prompt_len, prompts, firsts, vp,
                                               parity_tokens, page_size)
    n0 = ops.launch_count
    with ClockSampler(index=local) as clk:
        ms = _timed(lambda: (runner.ticks(K * P), runner.flush()))
    launches = torch.tensor([ops.launch_count - n0], device="cuda")
    dist.all_reduce(launches)

    # host-facing rounds: rank 0 (the stage the master addresses) copies every round's tokens of every slot to pinned
    # host memory and synchronises once per round, as a client streaming the tokens needs
    host_tok = torch.zeros(M, b, dtype=torch.int32).pin_memory()

    def host_rounds():
        for _ in range(K):
            runner.ticks(P)
            if rank == 0 and vp:
                for mb in range(M):
                    host_tok[mb].copy_(eng._state(b, mb)["tokens"], non_blocking=True)
                torch.cuda.current_stream().synchronize()
        runner.flush()

    stream_ms = _timed(host_rounds)
    for j in range(M):
        runner.close_slot(j)
    while runner.busy():
        runner.tick()
    runner.flush()
    torch.cuda.synchronize()

    S_mid = prompt_len + warm_rounds + K // 2
    stage_bytes = decode_step_bytes(spec, eng.n_local, eng.is_last and not vp, b, S_mid)
    if vp:
        stage_bytes += (eng.v1 - eng.v0) * spec.d_model * 2  # this stage's rows of the head
    tick_ms = ms / (K * P)
    busy = min(M, P) / P  # share of ticks in which a stage has a slot to work on
    ach = torch.tensor([stage_bytes / (tick_ms / 1e3) / 1e9], device="cuda")
    dist.all_reduce(ach, op=dist.ReduceOp.MIN)
    # what a pure layer pipeline can reach at this batch: every stage re-reads its weights once per slot
    ceiling = M * b / (P * (stage_bytes / (peaks["hbm_gbs"] * 1e9)))
    res = {"value": M * b * K / (ms / 1e3), "ms_per_step": ms / K, "gpu_launches": int(launches.item()),
           "clocks": clk.summary(), "tick_ms": tick_ms, "prefill_ms": pf_ms,
           "prefill_tokens_per_s": M * b * prompt_len / (pf_ms / 1e3), "prefill_chunks": stage.chunks,
           "stream_tokens_per_s": M * b * K / (stream_ms / 1e3), "stage_bytes": stage_bytes,
           "stage_pass_gbs": ach.item(), "stage_pass_frac": ach.item() / peaks["hbm_gbs"], "busy_ticks": busy,
           "pp_ceiling_tokens_per_s": ceiling, "slots": M, "b": b, "period": P, "hop": hop, "vp": vp,
           "prompt_len": prompt_len, "ctx_mid": S_mid, "parity_check": parity,
           "workload": f"{model_name} bf16 random-init, {world} pipeline stages (contiguous layer ranges, reference "
                       f"sharder formula) on {world} GPUs, {M} micro-batch slots x {b} sequences greedy decode at "
                       f"{S_mid}-token context, mailbox hand-off (one NCCL all-gather per tick, hop {hop})"}
    eng.unload()
    del eng, stage, ms_stage, runner
    torch.cuda.empty_cache()
    return res


def _parity_against_single_engine(spec, eng, rank, world, local, b, M, prompt_len, prompts, firsts, vp, n_tok, page_size):
    """Gather the pipeline's first n_tok tokens of every slot on rank 0 and check them against the single
    full-range engine holding the same (per-layer seeded) weights.  Returns "ok" / "ok (k ties)" / a skip reason;
    raises RuntimeError on a real mismatch (every rank raises, so torchrun exits non-zero)."""
    from .engine import ShardEngine

    dev = torch.device(f"cuda:{local}")
    holder = 0 if vp else world - 1
    toks = torch.zeros(M, n_tok, b, dtype=torch.int32, device=dev)
    if rank == holder:
        for mb in range(M):
            st = eng._state(b, mb)
            if vp:   # stage 0 logs the token it consumes at each step: log[0] is the prefill's token
                toks[mb] = st["token_log"][:n_tok]
            else:    # the last stage logs what it produces: the prefill's token, then log[0..]
                toks[mb, 0] = firsts[mb]
                toks[mb, 1:] = st["token_log"][:n_tok - 1]
    dist.broadcast(toks, src=holder)
    verdict = torch.zeros(2, dtype=torch.int32, device=dev)  # (status: 0 ok, 1 mismatch, 2 skipped; ties)
    total_w = spec.n_layers * spec.layer_weight_bytes() + 2 * spec.head_weight_bytes()
    if rank == 0:
        free = torch.cuda.mem_get_info()[0]
        if free < total_w + (12 << 30):
            verdict[0] = 2
        else:
            full = ShardEngine(spec, device=str(dev), max_batch=b, max_seq_len=prompt_len + n_tok + 8, page_size=page_size)
            full.load_random(seed=0)
            ties = 0
            for mb in range(M):
                pipe = toks[mb].t().contiguous()  # [b, n_tok]
                ids = prompts[mb].view(b, prompt_len).long()
                out, logits = full.generate(ids, prompt_len + n_tok, return_logits=True, forced=pipe[:, :-1])
                own = out[:, prompt_len:]
                for i in range(n_tok):
                    for r in range(b):
                        if int(own[r, i]) == int(pipe[r, i]):
                            continue
                        row = logits[i][r]
                        if float(row[int(pipe[r, i])]) >= float(row.max()) - 2.0 ** -7 * max(1.0, abs(float(row.max()))):
                            ties += 1  # the two tokens' logits tie to the last bf16 bit: either is a greedy choice
                        else:
                            verdict[0] = 1
            verdict[1] = ties
            full.unload()
            del full
            torch.cuda.empty_cache()
    dist.broadcast(verdict, src=0)
    status, ties = int(verdict[0]), int(verdict[1])
    if status == 1:
        raise RuntimeError("pipeline tokens differ from the single-GPU engine's on the same weights and prompts")
    if status == 2:
        return "skipped: the full model does not fit beside this stage on one GPU"
    return "ok" if ties == 0 else f"ok ({ties} bf16-level ties)"


_CTRL = {}


def serve_e2e(spec, model_name, rank, world, local, ctx, n_short, n_long, hop, page_size=16):
    """End to end through the plugin call at N GPUs: the worker group of pipeline_worker.py with this rank's stage
    injected (random-init weights, nothing read from disk), and hop*N CONCURRENT `POST /inference` requests on the
    head worker (werkzeug test client, no socket), each a `ctx`-token prompt given as text, greedy.  Every request's
    prompt ids travel host -> device, its tokens device -> host, inside the timed region.  Decode rate = the extra
    tokens of the long batch over the short one / the extra wall time; the whole batch's seconds are reported too."""
    import threading
    import time
    from concurrent.futures import ThreadPoolExecutor

    from .pipeline_worker import BLOCK, make_group_worker
    from .sharding import shard_metadata
    from .worker import IdTokenizer

    if "g" not in _CTRL:
        import datetime

        _CTRL["g"] = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=2))
    w = make_group_worker(rank, world, local, _CTRL["g"], env={"MODEL_CACHE_DIR": "/nonexistent", "USE_GPU": "1"}, hop=hop)
    P = w.backend.max_slots  # hop * (N + 1): the head is split over the stages
    eng = build_stage_engine(spec, rank, world, local, P, ctx + n_long + BLOCK + 8, page_size, vocab_parallel=True)
    eng.load_random(seed=0)
    name = "bench/" + model_name
    w.loaded_shards[name] = {rank: {"path": "(in memory)", "metadata": shard_metadata(name, rank, world, spec.n_layers)}}
    w.loaded_tokenizers[name] = IdTokenizer()
    w.loaded_models[f"{name}_{rank}"] = eng
    out = None
    if rank != 0:
        w.server.serve_forever()
    else:
        from werkzeug.test import Client

        loop = threading.Thread(target=lambda: (torch.cuda.set_device(local), w.server.serve_forever()))
        loop.start()
        try:
            g = torch.Generator().manual_seed(11)
            texts = [" ".join(str(int(t)) for t in torch.randint(0, spec.vocab, (ctx,), generator=g)) for _ in range(P)]

            def one(args):
                text, n_new = args
                r = Client(w).post("/inference", json={"model_name": name, "prompt": text, "shard_ids": list(range(world)),
                                                    "max_length": ctx + n_new, "do_sample": False, "timeout": 600})
                body = r.get_json()
                if r.status_code != 200 or len(body["result"].split()) != ctx + n_new:
                    raise RuntimeError(f"/inference on the worker group failed: {r.status_code} {str(body)[:200]}")

            def batch(n_new):
                t0 = time.perf_counter()
                with ThreadPoolExecutor(max_workers=P) as ex:
                    list(ex.map(one, [(t, n_new) for t in texts]))
                return time.perf_counter() - t0

            batch(n_short)  # warm-up: weights resident, graphs of every slot captured, P2P connections up
            t_short = batch(n_short)
            t_long = batch(n_long)
            extra = P * (n_long - n_short)
            steps = n_long - n_short
            out = {"value": extra / max(t_long - t_short, 1e-9), "unit": "tokens/s",
                   "h2d_bytes_per_step": 4.0 * ctx * P / steps, "d2h_bytes_per_step": 4.0 * P * n_long / steps,
                   "ms_per_step": (t_long - t_short) / steps * 1e3,
                   "api": f"{P} concurrent POST /inference on the worker group's head (werkzeug test client) -> "
                          "pipeline_worker.PipelineServer -> MailboxRunner",
                   "concurrent_requests": P, "request_batch_s": t_long, "new_tokens_per_request": n_long,
                   "request_tokens_per_s": P * n_long / t_long, "short_batch_s": t_short, "prompt_tokens": ctx,
                   "server_stats": dict(w.server.stats),
                   "note": "value = extra tokens of the long batch over the short one / extra wall time"}
        finally:
            w.server.shutdown()
            loop.join()
    eng.unload()
    torch.cuda.empty_cache()
    dist.barrier()
    return out


# BASELINE.json configs run on exactly N GPUs: (model, total batch, context, note)
NAMED_CONFIGS = {
    2: ("tinyllama-1.1b", 1, 1024, "configs[1]: TinyLlama-1.1B, 2 layer-range shards, batch 1 greedy decode"),
    4: ("mistral-7b", 64, 4096, "configs[3]: Mistral-7B, 4 shards, micro-batched decode, batch 64, 4k context"),
    8: ("llama-3-70b", 32, 8192, "configs[4]: Llama-3-70B, 8 shards, batch 32 decode at 8k context"),
}


def run_pipeline_bench(args, spec, peaks):
    """bench.py's N > 1 arm (launched by torchrun, one rank per GPU)."""
    rank, world, local = init_distributed()
    b, ctx, K, W = args.batch, args.ctx, args.steps, max(args.warmup, 3)
    vp = os.environ.get("DLI_PIPE_VP", "1") == "1"  # lm_head split by vocabulary rows over the stages
    hop = int(os.environ.get("DLI_PIPE_HOP", "2"))  # 2: the exchange runs under the next tick's kernels
    M = mailbox_period(1, world, hop, vp)           # slots in flight: every stage busy every tick
    main = measure_pipeline(spec, args.model, rank, world, local, b, M, ctx, K, W, args.page_size, peaks, vp=vp,
                            hop=hop, parity_tokens=8)
    named = None
    if world in NAMED_CONFIGS and not args.no_extra:
        from .engine import NAMED_SPECS

        name, B, nctx, note = NAMED_CONFIGS[world]
        try:
            # the stated batch split into one slot per stage (slots of B/N sequences; a single slot at batch 1);
            # hop 1: with no more slots than stages there is nothing to overlap the exchange with
            slots = min(world, B)
            nspec, nb = NAMED_SPECS[name], B // slots
            # Head layout by bytes per round of the slots: split over the stages, every stage streams its layers plus
            # V/N rows of the head on N+1 ticks (one tick of the period is idle with N slots); whole on the last
            # stage, that stage streams layers + the whole head on N ticks.  Mistral's 32k-row head is cheaper kept
            # whole (4 x 3.75 GB against 5 x 3.56 GB), the 128k-row heads are cheaper split.
            from .measure import decode_step_bytes
            from .sharding import shard_ranges as _ranges

            n_last = _ranges(nspec.n_layers, world)[-1]
            stage = decode_step_bytes(nspec, n_last[1] - n_last[0] + 1, False, nb, nctx)
            head = nspec.head_weight_bytes()
            named_vp = B > 1 and (world + 1) * (stage + head / world) < world * (stage + head)
            r = measure_pipeline(nspec, name, rank, world, local, nb, slots, nctx, min(K, 16), 3,
                                 args.page_size, peaks, vp=named_vp, hop=1, prompt_seed=3,
                                 parity_tokens=4 if name != "llama-3-70b" else 0, timed_prefill=False)
            named = {"config": note, "decode_tokens_per_s": r["value"], "batch": B, "ctx": r["ctx_mid"],
                     "micro_batches": r["slots"], "sequences_per_micro_batch": r["b"], "tick_ms": r["tick_ms"],
                     "head": "split by vocabulary rows over the stages" if named_vp else "whole, on the last stage",
                     "stage_pass_frac": r["stage_pass_frac"], "busy_ticks": r["busy_ticks"],
                     "pp_only_ceiling_tokens_per_s": r["pp_ceiling_tokens_per_s"],
                     "prefill_tokens_per_s": r["prefill_tokens_per_s"], "prefill_chunks": r["prefill_chunks"],
                     "parity_check": r["parity_check"], "clocks": r["clocks"], "workload": r["workload"],
                     "note": "a pure layer pipeline re-reads every stage's weights once per micro-batch, so its ceiling is "
                             "batch / (period x stage-pass time at the HBM peak), not the once-per-step figure of BASELINE.md 4"}
        except RuntimeError as e:
            if "pipeline tokens differ" in str(e):
                raise
            named = {"config": note, "error": str(e)[:300]}
    e2e = None
    try:
        e2e = serve_e2e(spec, args.model, rank, world, local, ctx, 8, 8 + max(16, min(K, 64)) // 4 * 4, hop, args.page_size)
    except Exception as e:  # noqa: BLE001 — the line still carries the device-timed numbers
        e2e_err = f"{type(e).__name__}: {str(e)[:200]}"
    roof = {"bound": "hbm", "kernel": "stage pass (all kernels of one micro-batch on one stage)",
            "achieved": main["stage_pass_gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": main["stage_pass_frac"],
            "traffic": None, "bytes_per_launch": main["stage_bytes"], "micro_batches": main["slots"],
            "note": "min over stages; every stage re-reads its weights once per micro-batch (SURVEY 8d)"}
    Mb = main["slots"] * main["b"]
    return {"value": main["value"], "ms_per_step": main["ms_per_step"], "gpu_launches": main["gpu_launches"],
            "clocks": main["clocks"],
            "e2e": e2e if e2e is not None else {
                "value": main["stream_tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 4 * Mb, "ms_per_step": Mb * 1e3 / main["stream_tokens_per_s"],
                "api": "MailboxRunner.ticks + per-round copy of the new tokens to pinned host memory on rank 0"
                       + ("" if rank != 0 else f" (the worker-group request path failed: {e2e_err})")},
            "roofline": roof,
            "extra": {"prefill_tokens_per_s": main["prefill_tokens_per_s"], "prefill_ms": main["prefill_ms"],
                      "prefill_chunks": main["prefill_chunks"], "tick_ms": main["tick_ms"], "period_ticks": main["period"],
                      "hop": hop, "parity_check": main["parity_check"], "peaks": peaks["source"],
                      "stream_tokens_per_s": main["stream_tokens_per_s"],
                      "pp_only_ceiling_tokens_per_s": main["pp_ceiling_tokens_per_s"], "named_config": named,
                      "env": {k: v for k, v in os.environ.items() if k.startswith("DLI_")}},
            "config": {"workload": main["workload"], "batch": Mb, "micro_batches": main["slots"], "ctx": main["ctx_mid"],
                       "page_size": args.page_size,
                       "parallelism": f"pp{world}" + ("+vocab-parallel head" if vp else ""),
                       "l2": "inputs larger than L2: each stage pass streams its weight slice "
                             f"({main['stage_bytes'] / 1e9:.2f} GB) from HBM"}}
