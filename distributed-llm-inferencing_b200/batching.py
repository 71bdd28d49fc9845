"""Continuous batching of concurrent /inference requests on one full-range ShardEngine.

The reference serves strictly one request at a time (one sync gunicorn worker,
worker/Dockerfile:45).  Decode at batch 1 is a pure weight
stream, so up to `max_batch` concurrent requests can share every weight read: each
request is prefilled alone into its own sequence slot (its own KV pages), then all
active slots decode together — one token per slot per step — and requests join and
leave between steps.  Batch rows are independent in every kernel, so a request gets
the same tokens whatever else is in flight (tests/test_batching_gpu.py).

One thread owns the engine (`serve_forever`, or `step()` driven by the caller); any
number of threads call `submit()` and wait on the returned Future.  Greedy decoding;
`max_length` counts prompt + new tokens like the reference's (worker/app.py:259).
"""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from .engine import check_max_length


@dataclass
class _Request:
    ids: torch.Tensor  # int64 [T] on the host
    max_length: int
    deadline: Optional[float]
    future: Future = field(default_factory=Future)
    out: List[int] = field(default_factory=list)
    do_sample: bool = False
    seed: int = 0


class ContinuousBatcher:
    def __init__(self, engine, max_batch: Optional[int] = None):
        if not (engine.is_first and engine.has_head):
            raise ValueError("continuous batching needs a full-range engine")
        self.eng = engine
        self.B = max_batch or engine.max_batch
        if self.B > engine.max_batch:
            raise ValueError("max_batch exceeds the engine's")
        self.pending: "queue.Queue[_Request]" = queue.Queue()
        self.slots: List[Optional[_Request]] = [None] * self.B
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.steps = 0
        self.grow = 64  # tokens of KV pages a sequence reserves ahead (the pool is shared on demand)
        st = engine._make_state(self.B, 0)  # the joint decode state: row i <-> sequence i
        st["seq_lens"].zero_()               # all rows idle
        st["positions"].zero_()
        self._tok_host = torch.zeros(self.B, dtype=torch.int32).pin_memory()
        # per-row token choice: greedy, or the reference's sampling call (temperature 0.8, top-k 50, top-p 0.95,
        # worker/app.py:297-305) with the request's own seed; the draw of the token at position p is keyed by
        # (seed, p), so a request gets the same tokens whichever row it lands in and whoever shares the batch
        self._params = dict(temperature=0.8, top_k=50, top_p=0.95)
        dev = engine.device
        self._rows = {"seeds": torch.zeros(self.B, dtype=torch.int64, device=dev), "steps": st["positions"],
                      "mask": torch.zeros(self.B, dtype=torch.int32, device=dev), "params": self._params}
        st["row_sampling"] = self._rows

    # ------------------------------------------------------------------ client side
    def submit(self, input_ids, max_length: int = 100, timeout: Optional[float] = None, do_sample: bool = False,
               seed: int = 0) -> Future:
        ids = torch.as_tensor(input_ids, dtype=torch.int64).reshape(-1).cpu()
        req = _Request(ids, int(max_length), None if timeout is None else time.time() + timeout,
                       do_sample=bool(do_sample), seed=int(seed))
        req.future.request = req  # a streaming caller reads req.out (the tokens so far) while it waits
        if len(ids) == 0:
            req.future.set_exception(ValueError("empty prompt"))
        elif req.max_length > self.eng.max_seq_len:
            req.future.set_exception(ValueError(f"max_length {req.max_length} > engine max_seq_len {self.eng.max_seq_len}"))
        elif req.max_length <= len(ids):  # transformers raises here; the worker answers 500 like the reference
            try:
                check_max_length(len(ids), req.max_length)
            except ValueError as e:
                req.future.set_exception(e)
        else:
            self.pending.put(req)
        return req.future

    # ------------------------------------------------------------------ engine-owner side
    def _admit(self) -> None:
        """Prefill waiting requests into free slots."""
        eng = self.eng
        joint = eng._state(self.B, 0)
        for i in range(self.B):
            if self.slots[i] is not None:
                continue
            try:
                req = self.pending.get_nowait()
            except queue.Empty:
                return
            T = len(req.ids)
            try:
                # sequence id i: pages for the prompt and the next GROW tokens; step() reserves more as it grows
                eng.start_batch(1, T, req.max_length, mb=i, reserve_len=T + self.grow)
            except RuntimeError as e:  # the shared page pool is empty right now
                if any(s is not None for s in self.slots):
                    self.pending.put(req)  # wait for a running request to leave
                else:
                    req.future.set_exception(e)  # nothing will ever free enough pages
                return
            ids = req.ids.to(device=eng.device, dtype=torch.int32)
            dev = eng.device
            self._rows["seeds"][i:i + 1].fill_(req.seed)
            self._rows["mask"][i:i + 1].fill_(int(req.do_sample))
            pst = eng._state(1, i)
            prefill_rows = {  # the prompt's last position keys the first new token
                "seeds": torch.tensor([req.seed], dtype=torch.int64, device=dev),
                "steps": torch.tensor([T - 1], dtype=torch.int32, device=dev),
                "mask": torch.tensor([int(req.do_sample)], dtype=torch.int32, device=dev), "params": self._params}
            # With max_batch == 1 the prefill state (1, 0) IS the joint decode state: its row_sampling (whose
            # tensors the captured decode graph reads) is only borrowed for the prefill and put back.
            keep = pst.get("row_sampling")
            pst["row_sampling"] = prefill_rows
            try:
                first = eng.prefill(1, T, ids=ids, mb=i)
            finally:
                if pst is joint:
                    pst["row_sampling"] = keep
            # hand the sequence over to row i of the joint decode state
            joint["tokens"][i:i + 1].copy_(first)
            joint["positions"][i:i + 1].fill_(T)
            joint["seq_lens"][i:i + 1].fill_(T + 1)
            req.out = [int(first.item())]
            self.slots[i] = req
            self._finish_if_done(i)

    def _finish_if_done(self, i: int) -> None:
        req = self.slots[i]
        # max_length reached, or the sequence produced an EOS id (transformers' stopping rule)
        done = len(req.ids) + len(req.out) >= req.max_length or (req.out and req.out[-1] in self.eng.eos_token_ids)
        expired = req.deadline is not None and time.time() > req.deadline
        if not (done or expired):
            return
        if expired and not done:
            req.future.set_exception(TimeoutError("Inference generation timed out"))
        else:
            req.future.set_result(torch.cat((req.ids, torch.tensor(req.out, dtype=torch.int64))))
        self.slots[i] = None
        self.eng._state(self.B, 0)["seq_lens"][i:i + 1].zero_()  # row idle again
        self.eng.kv.release(i)

    def step(self) -> int:
        """Admit what fits, then one decode step for every active slot.  Returns the number of
        active slots after the step."""
        self._admit()
        if not any(s is not None for s in self.slots):
            return 0
        eng = self.eng
        # paged KV on demand: a sequence about to outgrow its pages takes GROW tokens' worth more from the shared pool
        grew = False
        for i, req in enumerate(self.slots):
            if req is None:
                continue
            cur = len(req.ids) + len(req.out)  # tokens in the cache after this step
            if cur > eng.kv.capacity(i):
                try:
                    grew |= eng.kv.reserve(i, min(req.max_length, cur + self.grow))
                except RuntimeError as e:
                    req.future.set_exception(e)
                    self.slots[i] = None
                    eng._state(self.B, 0)["seq_lens"][i:i + 1].zero_()
                    eng.kv.release(i)
        if grew:
            eng.kv.commit()
        if not any(s is not None for s in self.slots):
            return 0
        toks = eng.decode_step(self.B)  # replayed CUDA graph over all rows (idle rows are no-ops for the caches)
        self._tok_host.copy_(toks, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.steps += 1
        for i, req in enumerate(self.slots):
            if req is not None:
                req.out.append(int(self._tok_host[i]))
                self._finish_if_done(i)
        return sum(s is not None for s in self.slots)

    def serve_forever(self, idle_sleep: float = 0.001) -> None:
        torch.cuda.set_device(self.eng.device)
        while not self._stop.is_set():
            try:
                if self.step() == 0 and self.pending.empty():
                    time.sleep(idle_sleep)
            except Exception as e:  # noqa: BLE001 — fail the requests in flight, keep serving
                for i, req in enumerate(self.slots):
                    if req is not None:
                        req.future.set_exception(e)
                        self.slots[i] = None

    def start(self) -> "ContinuousBatcher":
        self._thread = threading.Thread(target=self.serve_forever, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=10)
