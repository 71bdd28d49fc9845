"""N reference-style workers on the N GPUs of one box, their shards joined by an NCCL pipeline that serves
CONCURRENT requests.

The reference deploys one worker per node, registers shard i on node i (`/load_shard`, worker/app.py:139-206; one
node per (model, shard): master/dashboard/models.py:27-28) and sends `/inference` with the shards of the first node
that has any (master/dashboard/views.py:318-355) — whose worker then runs only its own shard (worker/app.py:336,
SURVEY.md A.3 Q2), one request at a time (one sync gunicorn worker, worker/Dockerfile:45).  Here

    python -m torch.distributed.run --nproc-per-node N -m dli_b200.pipeline_worker --port 5000

starts N workers in one process group: rank r is "node r" (HTTP on port+r, GPU r) with the same routes as a single
worker.  `/load_shard` registers shard r on rank r exactly as before.  An `/inference` that names shards of a model
split over the group (`metadata.json: num_shards == N`; the listed ids may be only the receiving node's, which is
what the reference master sends) is served by ALL ranks: a request arriving on another rank is handed to rank 0.

Serving loop (PipelineServer, one thread per rank, in lockstep):
  * the decode pipeline of pipeline.MailboxRunner runs continuously with one sequence per slot.  Llama-family shards
    in the sharder's directory layout get a vocabulary-parallel head (every stage reads its V/N rows of lm_head from
    the last shard's directory; hop * (N+1) slots): stage 0 takes the argmax over the stages' candidates, or — for a
    request that samples like the reference — draws from the whole logits row, whose slices travel in the mails,
    keyed by (the request's seed, the token's position).  Otherwise (GPT-2, head files elsewhere) the whole head
    stays on the last stage (hop * N slots).  All ranks agree on the layout per model;
  * every `block` rounds rank 0 reads the new tokens, completes the requests that reached `max_length`, an EOS id or
    their deadline, and broadcasts ONE plan over the control group (gloo): slots to close, slots to free, requests
    to admit.  Every rank applies the plan: a joining request is prefilled through the stages (hidden states by
    NCCL send/recv) into a free slot while the decode pipeline pauses between two ticks, then joins the ticks.
So a token of ONE request still costs a trip through the N stages, but up to hop * N requests share those trips —
what the reference's serialisation and the round-1 group (one request, one slot) could not do.
"""
from __future__ import annotations

import queue
import sys
import threading
import time
from concurrent.futures import Future
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

BLOCK = 4        # decode rounds between two plans
IDLE_BEAT = 20.0  # seconds between heartbeats of an idle group (keeps the control collective from timing out)


# ------------------------------------------------------------------------------ what a rank's GPU side must offer
class StageBackend:
    """The data plane of one rank as the serving loop sees it (CUDA: CudaBackend; the CPU tests plug in a toy)."""

    n_slots: int

    def probe(self, model: str) -> dict: return {}          # cheap facts about `model` here; every rank sees all of them
    def prepare(self, model: str, probes=None) -> None: ...  # make `model`'s stage resident (raise ValueError if it cannot be)
    def check_request(self, req: dict) -> None: ...          # raise ValueError if this rank cannot serve the request
    def join(self, slot: int, req: dict) -> None: ...        # prefill the request into `slot` (all ranks, collectively)
    def runner(self): ...                                    # the MailboxRunner of the resident model
    def read_tokens(self, slot: int, n: int) -> List[int]: ...  # rank 0: the first n tokens the slot has produced
    def release(self, slot: int) -> None: ...                # the slot's KV pages
    def reset(self) -> None: ...                             # forget the resident model (all ranks together)


class _Req:
    def __init__(self, cmd: dict):
        self.cmd = cmd
        self.future: Future = Future()
        self.slot: Optional[int] = None
        self.partial: List[int] = []  # the new tokens known so far (updated after every block)
        self.future.request = self


class PipelineServer:
    """Control plane.  Rank 0: submit() from any thread; every rank: serve_forever() in one thread."""

    def __init__(self, rank: int, world: int, ctrl_group, backend: StageBackend, block: int = BLOCK):
        self.rank, self.world, self.ctrl, self.be, self.block = rank, world, ctrl_group, backend, block
        self.queue: "queue.Queue[_Req]" = queue.Queue()
        self.model: Optional[str] = None
        self.active: Dict[int, _Req] = {}    # slot -> request (rank 0)
        self.closing: List[int] = []         # slots closed by the last plan: freed by the next one
        self.draining: List[int] = []
        self.rounds_done: Dict[int, int] = {}
        self._stop = threading.Event()
        self._failed = False                 # this rank's last block raised: told to the group with the next plan
        self._joining: List[_Req] = []
        self.stats = {"blocks": 0, "joined": 0, "max_active": 0}

    # ------------------------------------------------------------------ rank 0, any thread
    def submit(self, cmd: dict, on_partial=None) -> Future:
        """cmd: {model, ids, max_length, deadline, gen}.  The future yields the list of NEW token ids.
        on_partial(new_tokens_so_far) is called from the serving thread after every block (streaming)."""
        assert self.rank == 0
        r = _Req(cmd)
        r.on_partial = on_partial
        self.queue.put(r)
        return r.future

    def shutdown(self) -> None:
        self._stop.set()
        self.queue.put(None)

    # ------------------------------------------------------------------ rank 0, loop thread: the next plan
    def _plan(self) -> dict:
        """Blocks while there is nothing to do (heartbeats keep the other ranks' collective alive)."""
        last_beat = time.time()
        while True:
            if self._stop.is_set():
                return {"op": "stop"}
            joins, pend = [], []
            free = [s for s in range(self.be.n_slots) if s not in self.active and s not in self.closing
                    and s not in self.draining]
            while True:
                try:
                    r = self.queue.get_nowait()
                except queue.Empty:
                    break
                if r is None:
                    continue
                pend.append(r)
            keep = []
            for r in pend:
                model = r.cmd["model"]
                if self.model is not None and model != self.model and (self.active or joins or self.closing or self.draining):
                    keep.append(r)  # another model is being served: wait until its requests have left and their slots are freed
                elif not free:
                    keep.append(r)
                else:
                    if self.model is None or not (self.active or joins):
                        self.model = model  # (closing and draining are empty here: see the first branch)
                    r.slot = free.pop(0)
                    joins.append(r)
            for r in keep:
                self.queue.put(r)
            if joins or self.active or self.closing or self.draining:
                self._joining = joins
                return {"op": "run", "model": self.model, "close": list(self.closing), "free": list(self.draining),
                        "active": len(self.active),
                        "join": [(r.slot, {k: v for k, v in r.cmd.items() if k != "deadline"}) for r in joins],
                        "rounds": self.block}
            if time.time() - last_beat > IDLE_BEAT:
                return {"op": "idle"}
            try:
                r = self.queue.get(timeout=0.5)
                if r is not None:
                    self.queue.put(r)
            except queue.Empty:
                pass

    # ------------------------------------------------------------------ every rank
    def serve_forever(self) -> None:
        while True:
            box = [self._plan() if self.rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=self.ctrl)
            plan = box[0]
            if plan["op"] == "stop":
                return
            if plan["op"] == "idle":
                continue
            try:
                self._apply(plan)
            except Exception as e:  # noqa: BLE001 — a data-plane failure: fail what is in flight, keep serving
                print(f"[dli pipeline worker {self.rank}] block failed: {type(e).__name__}: {e}", file=sys.stderr, flush=True)
                # The slots' state (open rounds, KV pages) is unknown now.  Ranks must rebuild their stages TOGETHER (the
                # runners' tick counters address the mailboxes), so the failure is only noted here; the next plan's
                # first exchange tells every rank, and all of them start again from an empty pipeline.  (A failure every
                # rank meets at the same point — the KV pool running out in a join — or one on rank 0's bookkeeping is
                # repaired this way; a rank failing alone inside an NCCL call cannot be repaired from here.)
                self._failed = True
                if self.rank == 0:
                    for r in list(self.active.values()) + list(getattr(self, "_joining", [])):
                        if not r.future.done():
                            r.future.set_exception(e)
                    self.active.clear()
                    self._joining = []

    def _agree(self, errs: dict) -> dict:
        """Union of every rank's {key: message} (first rank wins per key), identical on all ranks."""
        box: List[Optional[dict]] = [None] * self.world
        dist.all_gather_object(box, errs, group=self.ctrl)
        out: dict = {}
        for d in box:
            for k, v in (d or {}).items():
                out.setdefault(k, v)
        return out

    def _apply(self, plan: dict) -> None:
        be = self.be

        def msg(e):
            return str(e) if isinstance(e, ValueError) else f"{type(e).__name__}: {e}"

        # 1a. what every rank knows about the model (is its shard registered, can the head be split): all ranks see
        #     all answers, so they take the same decisions ("*": a failure that concerns every joining request)
        errs, probe = {}, None
        try:
            probe = be.probe(plan["model"])
        except Exception as e:  # noqa: BLE001
            errs["*"] = msg(e)
        box: List[Optional[dict]] = [None] * self.world
        dist.all_gather_object(box, {"errs": errs, "probe": probe, "failed": self._failed}, group=self.ctrl)
        self._failed = False
        errs = {}
        for d in box:
            for k, v in d["errs"].items():
                errs.setdefault(k, v)
        failed = any(d["failed"] for d in box)
        if failed or (not errs and any((d["probe"] or {}).get("stale") for d in box)):
            # some rank's last block failed, or its resident engine was unloaded behind the pipeline's back: every rank
            # rebuilds its stage now (together: the runners' tick counters must agree); what was in flight is lost
            be.reset()
            if self.rank == 0:
                why = "a block of the pipeline failed" if failed else "the model was unloaded while the request was running"
                for r in self.active.values():
                    if not r.future.done():
                        r.future.set_exception(RuntimeError(why))
                self.active.clear()
            self.closing, self.draining = [], []
            plan = dict(plan, close=[], free=[], active=0)
        # 1b. the model's stage on every rank (may load weights), then every joining request on its own (a request
        #     one rank must refuse fails alone, for everybody)
        errs2: dict = {}
        if not errs:
            try:
                be.prepare(plan["model"], [d["probe"] for d in box])
            except Exception as e:  # noqa: BLE001
                errs2["*"] = msg(e)
            if not errs2:
                for slot, req in plan["join"]:
                    try:
                        if slot >= be.n_slots:
                            raise ValueError("no free slot")  # planned for a model with more slots: rank 0 re-queues it
                        be.check_request(req)
                    except Exception as e:  # noqa: BLE001
                        errs2[slot] = msg(e)
            errs = self._agree(errs2)
        joins = [(slot, req) for slot, req in plan["join"] if "*" not in errs and slot not in errs]
        if self.rank == 0:
            for r in self._joining:
                if "*" not in errs and errs.get(r.slot) == "no free slot":
                    r.slot = None
                    self.queue.put(r)
                elif "*" in errs or r.slot in errs:
                    r.future.set_exception(ValueError(errs.get("*", errs.get(r.slot))))
            self._joining = [r for r in self._joining if "*" not in errs and r.slot is not None and r.slot not in errs]
        if "*" in errs:
            # A model change that one rank could not follow (models only change while no request is active).  Ranks
            # that did switch and ranks that did not must not tick together — their runners' tick counters, which
            # address the mailboxes, would differ — so EVERY rank forgets its resident model; the next plan rebuilds
            # it everywhere at once.
            be.reset()
            self.model, self.closing, self.draining = None, [], []
            return
        runner = be.runner()
        # 2. slots whose last rounds have drained: their KV pages go back; slots to close: no further round starts
        for s in plan["free"]:
            be.release(s)
        for s in plan["close"]:
            runner.close_slot(s)
        # 3. admit: prompt pass through the stages, then the slot joins the ticks
        for slot, req in joins:
            be.join(slot, req)
            runner.open_slot(slot)
        # 4. the block
        runner.ticks(plan["rounds"] * runner.P)
        runner.flush()
        # 5. rank 0: bookkeeping for the next plan
        if self.rank == 0:
            self._after_block(plan, joins)

    def _after_block(self, plan: dict, joins) -> None:
        self.draining = list(plan["close"])  # closed in this block: their in-flight rounds are through by the next
        self.closing = []
        for r in getattr(self, "_joining", []):
            if r.slot is not None and any(s == r.slot for s, _ in joins):
                self.active[r.slot] = r
                self.rounds_done[r.slot] = 0
                self.stats["joined"] += 1
        self._joining = []
        self.stats["blocks"] += 1
        self.stats["max_active"] = max(self.stats["max_active"], len(self.active))
        now = time.time()
        for slot, r in list(self.active.items()):
            self.rounds_done[slot] += plan["rounds"]
            n_target = r.cmd["max_length"] - len(r.cmd["ids"])
            have = min(self.rounds_done[slot], n_target)  # tokens the slot has CONSUMED = produced and fed back
            toks = self.be.read_tokens(slot, have)
            r.partial = toks
            if getattr(r, "on_partial", None) is not None and toks:
                try:
                    r.on_partial(toks)
                except Exception:  # noqa: BLE001 — a client's progress sink must not stop the pipeline
                    pass
            eos = tuple(r.cmd.get("eos") or ())
            hit = next((i for i, t in enumerate(toks) if t in eos), None)
            if hit is not None:
                r.future.set_result(toks[: hit + 1])
            elif have >= n_target:
                r.future.set_result(toks[:n_target])
            elif r.cmd.get("deadline") is not None and now > r.cmd["deadline"]:
                r.future.set_exception(TimeoutError("Inference generation timed out"))
            else:
                continue
            del self.active[slot]
            self.closing.append(slot)


# ------------------------------------------------------------------------------ CUDA data plane of one rank
class CudaBackend(StageBackend):
    """This rank's ShardEngine of the resident model, in mailbox mode with one sequence per slot.

    Two head layouts, agreed by all ranks per model (probe -> prepare):
      * vocabulary-parallel (Llama-family shards in the sharder's directory layout): every stage holds V/N rows of the
        lm_head, read from the LAST shard's directory, so no stage carries the whole [V, d] matrix on top of its layers;
        stage 0 takes the argmax over the stages' candidates, or — for a request that samples — draws from the whole
        logits row, whose slices travel in the mails; hop*(N+1) slots;
      * the whole head on the last stage (GPT-2, or head files not found): hop*N slots."""

    def __init__(self, worker, rank: int, world: int, hop: int = 2):
        import os

        from .pipeline import mailbox_period

        self.worker, self.rank, self.world, self.hop = worker, rank, world, hop
        self.allow_vp = os.environ.get("DLI_GROUP_VP", "1") == "1"
        self.max_slots = mailbox_period(1, world, hop, True)
        self.n_slots = mailbox_period(1, world, hop, False)
        self.model = None
        self.eng = None
        self.vp = False
        self._runner = None
        self._stage = None
        self._rows = {}

    # -- which shard is mine, and can its head be split? ---------------------------------------------------------
    def _my_shard(self, model: str):
        shards = self.worker.loaded_shards.get(model, {})
        mine = [sid for sid, info in shards.items()
                if info["metadata"].get("num_shards") == self.world and info["metadata"].get("shard_id", sid) == self.rank]
        if not mine:  # registered under another id: matched by the group size alone
            mine = [sid for sid, info in shards.items() if info["metadata"].get("num_shards") == self.world]
        if not mine:
            raise ValueError(f"Shard {self.rank} of model {model} is not loaded")
        return mine[0], shards[mine[0]]

    def _head_dir(self, info):
        """Directory that stores the final norm and lm_head: the last shard's (this one's on the last rank; else its
        sibling `shard_<N-1>` in the sharder's layout, shard_model.py:94)."""
        import os

        if self.rank == self.world - 1:
            return info["path"]
        cand = os.path.join(os.path.dirname(os.path.normpath(info["path"])), f"shard_{self.world - 1}")
        return cand if os.path.exists(os.path.join(cand, "config.json")) else None

    def probe(self, model: str) -> dict:
        import json
        import os

        with self.worker._lock:
            sid, info = self._my_shard(model)
            eng = self.worker.loaded_models.get(f"{model}_{sid}")
        # stale: /unload_model (or a re-registration) took this rank's resident engine away since the last plan
        stale = self.model == model and self.eng is not None and eng is not self.eng
        if eng is not None:  # already resident (or put there by a host that built it itself): its layout decides
            return {"vp": eng.head_shard is not None, "stale": stale}
        cfg_path = os.path.join(info["path"], "config.json")
        arch = json.load(open(cfg_path)).get("model_type") if os.path.exists(cfg_path) else None
        return {"vp": bool(self.allow_vp and arch in ("llama", "mistral") and self._head_dir(info)), "stale": stale}

    def prepare(self, model: str, probes=None) -> None:
        if model == self.model and self.eng is not None:
            return
        from .pipeline import DistExchange, MailboxRunner, MailStage

        vp = bool(probes) and all(p and p.get("vp") for p in probes)
        w = self.worker
        with w._lock:
            sid, info = self._my_shard(model)
            md = info["metadata"]
            key = f"{model}_{sid}"
            if key not in w.loaded_models:  # lazy weight load on first use, like worker/app.py:345-349
                w.loaded_models[key] = w._engine(info["path"], md["start_layer"], md["end_layer"], max_batch=self.max_slots,
                                                 head_shard=(self.rank, self.world) if vp else None,
                                                 head_path=self._head_dir(info) if vp else None)
            eng = w.loaded_models[key]
        if (eng.head_shard is not None) != vp:
            raise ValueError(f"shard {sid} of model {model} is resident with another head layout than the group agreed on")
        if eng.is_first != (self.rank == 0) or eng.is_last != (self.rank == self.world - 1):
            raise ValueError(f"shard {sid} (layers {md['start_layer']}-{md['end_layer']}) is registered on GPU {self.rank}: "
                             "shard i of N must live on GPU i")
        n_slots = self.hop * (self.world + 1 if vp else self.world)
        if eng.max_batch < n_slots:
            raise ValueError(f"engine of shard {sid} was created for {eng.max_batch} sequences, the group needs {n_slots}")
        self.eng, self.model, self.vp, self.n_slots = eng, model, vp, n_slots
        self._stage = MailStage(eng, 1, n_slots, n_slots, self.hop, self.rank, self.world, with_logits=vp)
        comm = torch.cuda.Stream(device=eng.device) if self.hop > 1 else None
        self._runner = MailboxRunner(self._stage, self.rank, self.world, n_slots, hop=self.hop, vp=vp,
                                     exchange=DistExchange(comm_stream=comm))
        self._rows = {}

    def runner(self):
        return self._runner

    def reset(self) -> None:
        if self.eng is not None:
            self.eng.kv.release_all()  # the slots' pages: the engine may stay resident in the worker's registry
        self.model, self.eng, self._runner, self._stage, self._rows = None, None, None, None, {}

    def check_request(self, req: dict) -> None:
        from .engine import check_max_length

        T = len(req["ids"])
        check_max_length(T, req["max_length"])
        if req["max_length"] + BLOCK + 2 > self.eng.max_seq_len:
            raise ValueError(f"max_length {req['max_length']} exceeds what this worker group holds per sequence "
                             f"({self.eng.max_seq_len - BLOCK - 2}; DLI_MAX_SEQ_LEN)")

    def _slot_rows(self, slot: int) -> dict:
        """Persistent per-slot token-choice state on the stage that chooses (the last stage; stage 0 with the head split):
        its addresses are baked into the slot's graphs.  Greedy, or the reference's sampling call keyed by (the request's
        seed, the token's position)."""
        if slot not in self._rows:
            st, dev = self.eng._state(1, slot), self.eng.device
            steps = torch.zeros(1, dtype=torch.int32, device=dev) if self.vp else st["positions"]
            self._rows[slot] = {"seeds": torch.zeros(1, dtype=torch.int64, device=dev), "steps": steps,
                                "mask": torch.zeros(1, dtype=torch.int32, device=dev),
                                "params": dict(temperature=0.8, top_k=50, top_p=0.95)}
        return self._rows[slot]

    def join(self, slot: int, req: dict) -> None:
        from . import ops

        eng, rank, world = self.eng, self.rank, self.world
        T = len(req["ids"])
        eng.start_batch(1, T, req["max_length"] + BLOCK + 2, mb=slot)
        st = eng._state(1, slot)
        gen = req.get("gen") or {}
        sample = bool(gen.get("do_sample"))
        chooser = eng.is_first if self.vp else eng.has_head
        if chooser:
            rows = self._slot_rows(slot)
            rows["seeds"].fill_(int(gen.get("seed", 0)))
            rows["mask"].fill_(1 if sample else 0)
            # the prompt's last position keys the first new token; decode steps then key by the slot's position
            st["row_sampling"] = dict(rows, steps=torch.tensor([T - 1], dtype=torch.int32, device=eng.device))
        (self._stage.sampling.add if sample else self._stage.sampling.discard)(slot)
        d = eng.spec.d_model
        if rank == 0:
            ids = torch.tensor(req["ids"], dtype=torch.int32, device=eng.device)
            x = eng.prefill(1, T, ids=ids, mb=slot)
        else:
            buf = torch.empty(T, d, dtype=torch.bfloat16, device=eng.device)
            dist.recv(buf, src=rank - 1)
            x = eng.prefill(1, T, hidden=buf, mb=slot)
        if rank < world - 1:
            dist.send(x.contiguous(), dst=rank + 1)
        if not self.vp:
            if eng.has_head:
                dist.send(st["tokens"], dst=0)  # the first token goes to the pipeline head
            if rank == 0:
                dist.recv(st["tokens"], src=world - 1)
        else:
            # the last stage's final hidden state to every stage; each computes its head slice; stage 0 chooses
            dist.broadcast(st["hf"], src=world - 1)
            eng.head_slice_step(1, slot)
            cands = torch.empty(world, 1, 2, dtype=torch.float32, device=eng.device)
            dist.all_gather_into_tensor(cands.view(-1), st["cand"].reshape(-1).clone())
            if rank == 0:
                ops.argmax_merge(cands, st["tokens"])
            if sample:
                per = (eng.spec.vocab + world - 1) // world
                mine = torch.zeros(per, dtype=torch.bfloat16, device=eng.device)
                mine[: eng.v1 - eng.v0].copy_(st["logits"][0])
                row = torch.empty(world * per, dtype=torch.bfloat16, device=eng.device)
                dist.all_gather_into_tensor(row, mine)
                if rank == 0:
                    r = st["row_sampling"]
                    ops.sample_rows(row.view(1, -1)[:, : eng.spec.vocab], st["tokens"], r["seeds"], r["steps"], r["mask"],
                                    **r["params"])
        if chooser:
            st["row_sampling"] = self._slot_rows(slot)
        st["step"].zero_()

    def read_tokens(self, slot: int, n: int) -> List[int]:
        torch.cuda.current_stream().synchronize()
        return self.eng._state(1, slot)["token_log"][:n, 0].tolist()

    def release(self, slot: int) -> None:
        self.eng.kv.release(slot)
        self.eng._state(1, slot)["seq_lens"].zero_()  # an idle row: its decode steps touch nothing
        if self._stage is not None:
            self._stage.sampling.discard(slot)


# ------------------------------------------------------------------------------ the worker of one rank
def make_group_worker(rank: int, world: int, local: int, ctrl_group, env=None, base_port: int = 5000,
                      host: str = "127.0.0.1", hop: int = 2):
    """A Worker (same routes as worker.py) bound to GPU `local`, whose sharded /inference goes to the group's
    serving loop when the model is split over the group."""
    from .engine import apply_eos, check_max_length
    from .worker import Worker

    class GroupWorker(Worker):
        def __init__(self):
            super().__init__(env)
            self.cuda_device = f"cuda:{local}"
            self.backend = CudaBackend(self, rank, world, hop=hop)
            self.server = PipelineServer(rank, world, ctrl_group, self.backend)

        def _release_engine(self, key, eng) -> None:
            # The stage that is resident in the serving pipeline may be in the middle of a block of ticks on another
            # thread: only drop the references here.  The pipeline notices at its next plan (probe: "stale"), all
            # ranks rebuild together, and the memory goes when the old runner and its graphs are released.
            if eng is self.backend.eng:
                return
            super()._release_engine(key, eng)

        def dispatch(self, method, path, data, headers):
            if torch.cuda.is_available():
                torch.cuda.set_device(local)  # HTTP handler threads start on device 0; this worker owns GPU `local`
            if method == "POST" and path == "/inference" and rank != 0 and self._is_group_request(data or {}):
                return self._forward_to_head(data, headers)
            return super().dispatch(method, path, data, headers)

        def _is_group_request(self, data) -> bool:
            """Does the request name shards of a model that is split over the whole group (num_shards == N)?"""
            # whichever ids the request lists (the reference master lists only the ids of the node it picked)
            shards = self.loaded_shards.get(data.get("model_name"), {})
            return world > 1 and bool(data.get("shard_ids")) and any(
                info["metadata"].get("num_shards") == world for info in shards.values())

        def _forward_to_head(self, data, headers):
            """The reference master posts to whichever node it finds first (views.py:337-340): hand the request to
            rank 0, the pipeline head, and relay its answer."""
            import requests

            hdr = {"Authorization": headers["Authorization"]} if headers.get("Authorization") else {}
            r = requests.post(f"http://{host}:{base_port}/inference", json=data, headers=hdr,
                              timeout=float(data.get("timeout", 60)) + 60)
            return r.json(), r.status_code

        def run_shard_inference(self, model_name, prompt, shard_ids, max_length, deadline=None, gen=None):
            local_shards = self.loaded_shards.get(model_name, {})
            group_model = world > 1 and any(i["metadata"].get("num_shards") == world for i in local_shards.values())
            if not group_model:
                if world > 1 and not all(s in local_shards for s in shard_ids):
                    raise ValueError(f"Shard {[s for s in shard_ids if s not in local_shards][0]} of model {model_name} is not loaded")
                return super().run_shard_inference(model_name, prompt, shard_ids, max_length, deadline, gen)
            tokenizer = self.loaded_tokenizers[model_name]
            ids = tokenizer(prompt, return_tensors="pt")["input_ids"][0]
            check_max_length(len(ids), max_length)
            eos, pad = self._stop_ids(model_name)
            on = self._on_tokens(tokenizer)  # streaming: the text so far after every block of rounds
            n_new = int(max_length) - len(ids)
            fut = self.server.submit(
                {"model": model_name, "ids": ids.tolist(), "max_length": int(max_length), "deadline": deadline,
                 "gen": gen or {}, "eos": list(eos)},
                on_partial=None if on is None else (
                    lambda toks: on(torch.cat((ids.long(), torch.tensor(toks[:n_new], dtype=torch.int64)))[None])))
            wait = None if deadline is None else max(1.0, deadline - time.time() + 30)
            new = fut.result(timeout=wait)
            new = apply_eos(torch.tensor([new], dtype=torch.int64), eos, pad)
            return tokenizer.decode(torch.cat((ids.long(), new[0])), skip_special_tokens=True)

        def _stop_ids(self, model_name):
            import json
            import os

            from .worker import _stop_tokens

            for info in self.loaded_shards.get(model_name, {}).values():
                p = info["path"]
                if os.path.exists(os.path.join(p, "config.json")):
                    return _stop_tokens(p, json.load(open(os.path.join(p, "config.json"))))
            return (), None

        def inference(self, data):
            # the group's requests wait on their future OUTSIDE the worker lock, so that they overlap
            data = data or {}
            if (rank == 0 and data.get("shard_ids") and data.get("model_name") and data.get("prompt")
                    and self._is_group_request(data)):
                start = time.time()
                try:
                    result = self.run_shard_inference(data["model_name"], data["prompt"], data["shard_ids"],
                                                      data.get("max_length", 100), start + data.get("timeout", 60),
                                                      self._gen_kwargs(data))
                    return {"status": "success", "result": result, "execution_time": time.time() - start}, 200
                except TimeoutError as e:
                    return {"status": "error", "message": str(e) or "Inference generation timed out"}, 408
                except Exception as e:  # noqa: BLE001
                    return {"status": "error", "message": f"Inference failed: {str(e)}"}, 500
            return super().inference(data)

    return GroupWorker()


def main(argv=None):
    import argparse
    import datetime
    import os

    from werkzeug.serving import make_server

    from .pipeline import init_distributed

    ap = argparse.ArgumentParser(description="N workers on N GPUs, shards joined by an NCCL pipeline; rank r "
                                             "serves the reference worker's routes on port + r")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=5000)
    a = ap.parse_args(argv)
    rank, world, local = init_distributed()
    ctrl = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=12))
    head_host = "127.0.0.1" if a.host in ("0.0.0.0", "") else a.host
    w = make_group_worker(rank, world, local, ctrl, base_port=a.port, host=head_host,
                          hop=int(os.environ.get("DLI_PIPE_HOP", "2")))
    srv = make_server(a.host, a.port + rank, w, threaded=True)
    http = threading.Thread(target=srv.serve_forever, daemon=True)
    http.start()
    torch.cuda.set_device(local)
    try:
        w.server.serve_forever()  # every rank: the serving loop (rank 0 plans, the others follow)
    finally:
        srv.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
