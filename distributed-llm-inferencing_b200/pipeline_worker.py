"""N reference-style workers on the N GPUs of one box, their shards joined by an NCCL pipeline.

The reference deploys one worker per node, registers shard i on node i (`/load_shard`,
worker/app.py:139-206; one node per (model, shard): master/dashboard/models.py:27-28) and sends
`/inference` with the shard list to the first node (master/dashboard/views.py:318-355) — whose
worker then runs only its own shard (worker/app.py:336).  Here

    python -m torch.distributed.run --nproc-per-node N -m dli_b200.pipeline_worker --port 5000

starts N workers in one process group: rank r is "node r" (HTTP on port+r, GPU r) with the same
routes as a single worker.  `/load_shard` registers shard r on rank r exactly as before.  When rank
0 receives an `/inference` whose `shard_ids` name shards that live on the other ranks, it tells
them over a control group (gloo) and all ranks run the layer-range pipeline of pipeline.py: hidden
states travel GPU to GPU by NCCL send/recv, the last stage chooses the tokens (greedy or the
reference's sampling call) and they come back to rank 0, which answers the request.  Shard i must
be registered on rank i (shard i <-> GPU i).

The generation runs in blocks of a few decode rounds; between blocks the LAST stage (it sees the
tokens) decides whether every sequence has hit an EOS id or the deadline has passed and broadcasts
that one flag, so all ranks leave the loop together.
"""
from __future__ import annotations

import sys
import time
from typing import Callable, Dict, List, Optional

import torch
import torch.distributed as dist

BLOCK = 8  # decode rounds between two stop checks


class PipelineGroup:
    """Control plane of the worker group.  `stage_factory(cmd)` returns the object that runs this
    rank's part of a request (see CudaStageRun below; the CPU tests plug in a toy)."""

    def __init__(self, rank: int, world: int, ctrl_group, stage_factory: Callable[[dict], "StageRun"]):
        self.rank, self.world, self.ctrl, self.factory = rank, world, ctrl_group, stage_factory

    # ------------------------------------------------------------------ rank 0: called by the HTTP handler
    def generate(self, cmd: dict) -> List[int]:
        """Broadcast the request and take part in it; returns the new tokens (rank 0 only)."""
        assert self.rank == 0
        dist.broadcast_object_list([dict(cmd, op="generate")], src=0, group=self.ctrl)
        return self._run(cmd)

    def shutdown(self) -> None:
        assert self.rank == 0
        dist.broadcast_object_list([{"op": "stop"}], src=0, group=self.ctrl)

    # ------------------------------------------------------------------ ranks > 0: main loop
    def serve_forever(self) -> None:
        while True:
            box = [None]
            dist.broadcast_object_list(box, src=0, group=self.ctrl)
            cmd = box[0]
            if cmd.get("op") == "stop":
                return
            try:
                self._run(cmd)
            except Exception as e:  # noqa: BLE001 — rank 0 raises the same error to the client (see _run)
                print(f"[dli pipeline worker {self.rank}] request failed: {type(e).__name__}: {e}", file=sys.stderr,
                      flush=True)

    # ------------------------------------------------------------------ every rank
    def _run(self, cmd: dict) -> Optional[List[int]]:
        # 1. every rank builds (or finds) its stage; any failure is agreed on before the pipeline starts
        run, err = None, None
        try:
            run = self.factory(cmd)
        except Exception as e:  # noqa: BLE001
            err = f"{type(e).__name__}: {e}" if not isinstance(e, ValueError) else str(e)
        errs: List[Optional[str]] = [None] * self.world
        dist.all_gather_object(errs, err, group=self.ctrl)
        bad = [e for e in errs if e]
        if bad:
            raise ValueError(bad[0])
        # 2. prefill, then decode in blocks; the last stage decides when to stop
        n_new = cmd["max_length"] - len(cmd["ids"])
        run.prefill()
        done = 1  # tokens produced so far
        reason = 0
        while done < n_new:
            k = min(BLOCK, n_new - done)
            run.decode(k)
            done += k
            reason = run.agree_stop(cmd.get("deadline"))
            if reason:
                break
        # 3. the last stage owns the token log; rank 0 answers the request
        box = [run.tokens(done) if self.rank == self.world - 1 else None]
        dist.broadcast_object_list(box, src=self.world - 1, group=self.ctrl)
        run.close()
        if reason == 2:
            raise TimeoutError("Inference generation timed out")
        return box[0] if self.rank == 0 else None


class StageRun:
    """What PipelineGroup needs from a rank's part of one request."""

    def prefill(self) -> None: ...
    def decode(self, rounds: int) -> None: ...
    def agree_stop(self, deadline: Optional[float]) -> int: ...  # 0 = go on, 1 = every sequence finished, 2 = deadline
    def tokens(self, n: int) -> List[int]: ...  # last stage: the n new tokens
    def close(self) -> None: ...


# ------------------------------------------------------------------------------ CUDA stage of one request
class CudaStageRun(StageRun):
    """This rank's ShardEngine driven by pipeline.PipelineRunner for one request (one slot, batch 1)."""

    def __init__(self, worker, rank: int, world: int, cmd: dict):
        from .engine import all_finished  # noqa: F401  (used in agree_stop)
        from .pipeline import EngineStage, PipelineRunner

        model, shard_ids = cmd["model"], list(cmd["shard_ids"])
        if len(shard_ids) != world:
            raise ValueError(f"{len(shard_ids)} shards requested but the worker group has {world} GPUs "
                             "(one shard per GPU)")
        sid = sorted(shard_ids)[rank]  # shard i <-> GPU i
        with worker._lock:
            if sid not in worker.loaded_shards.get(model, {}):
                raise ValueError(f"Shard {sid} of model {model} is not loaded")
            info = worker.loaded_shards[model][sid]
            md = info["metadata"]
            key = f"{model}_{sid}"
            if key not in worker.loaded_models:  # lazy weight load on first use, like worker/app.py:345-349
                worker.loaded_models[key] = worker._engine(info["path"], md["start_layer"], md["end_layer"])
            eng = worker.loaded_models[key]
        if eng.is_first != (rank == 0) or eng.is_last != (rank == world - 1):
            raise ValueError(f"shard {sid} (layers {md['start_layer']}-{md['end_layer']}) is registered on GPU {rank}: "
                             "shard i of N must live on GPU i")
        self.eng, self.rank, self.world = eng, rank, world
        ids = torch.tensor(cmd["ids"], dtype=torch.int32, device=eng.device)
        T = len(cmd["ids"])
        self.stage = EngineStage(eng, 1, T, cmd["max_length"], 1)
        if eng.is_first:
            self.stage.prompt_ids[0] = ids
        gen = cmd.get("gen") or {}
        mode = (dict(temperature=float(gen.get("temperature", 0.8)), top_k=int(gen.get("top_k", 50)),
                     top_p=float(gen.get("top_p", 0.95))) if gen.get("do_sample") else None)
        if mode != eng.sampling:  # the token choice lives on the last stage; harmless elsewhere
            eng.sampling = mode
            eng._graphs.clear()
        st = eng._state(1, 0)
        st.pop("row_sampling", None)
        st["seed"].fill_(int(gen.get("seed", 0)))
        st["step"].fill_(-1)
        self.runner = PipelineRunner(self.stage, rank, world, 1, return_last=True)
        self.flag = torch.zeros(1, dtype=torch.int32, device=eng.device)
        self.first = None
        self.done = 0

    def prefill(self) -> None:
        self.runner.run("prefill")
        st = self.eng._state(1, 0)
        if self.eng.is_last:
            self.first = st["tokens"].clone()
        st["step"].zero_()
        self.done = 1

    def decode(self, rounds: int) -> None:
        self.runner.run("decode", rounds)
        self.done += rounds

    def _seen(self) -> torch.Tensor:  # last stage: [done, 1] tokens so far
        log = self.eng._state(1, 0)["token_log"]
        return torch.cat((self.first[None], log[: self.done - 1]), dim=0)

    def agree_stop(self, deadline: Optional[float]) -> int:
        from .engine import all_finished

        if self.rank == self.world - 1:
            torch.cuda.current_stream().synchronize()
            reason = 0
            if all_finished(self._seen(), self.eng.eos_token_ids):
                reason = 1
            elif deadline is not None and time.time() > deadline:
                reason = 2
            self.flag.fill_(reason)
        dist.broadcast(self.flag, src=self.world - 1)
        return int(self.flag.item())

    def tokens(self, n: int) -> List[int]:
        return self._seen()[:n, 0].tolist()

    def close(self) -> None:
        pass


# ------------------------------------------------------------------------------ the worker of one rank
def make_group_worker(rank: int, world: int, local: int, ctrl_group, env=None):
    """A Worker (same routes as worker.py) bound to GPU `local`, whose sharded /inference runs the
    NCCL pipeline when the requested shards are not all resident on this rank."""
    from .engine import apply_eos, check_max_length
    from .worker import Worker

    class GroupWorker(Worker):
        def __init__(self):
            super().__init__(env)
            self.cuda_device = f"cuda:{local}"
            self.group = PipelineGroup(rank, world, ctrl_group, lambda cmd: CudaStageRun(self, rank, world, cmd))

        def dispatch(self, method, path, data, headers):
            if torch.cuda.is_available():
                torch.cuda.set_device(local)  # HTTP handler threads start on device 0; this worker owns GPU `local`
            return super().dispatch(method, path, data, headers)

        def run_shard_inference(self, model_name, prompt, shard_ids, max_length, deadline=None, gen=None):
            local_shards = self.loaded_shards.get(model_name, {})
            if world == 1 or all(s in local_shards for s in shard_ids):
                return super().run_shard_inference(model_name, prompt, shard_ids, max_length, deadline, gen)
            if rank != 0:
                raise ValueError("send /inference for a model sharded over the GPUs to the first worker (rank 0)")
            if model_name not in self.loaded_tokenizers:
                raise ValueError(f"Shard {sorted(shard_ids)[0]} of model {model_name} is not loaded")
            tokenizer = self.loaded_tokenizers[model_name]
            ids = tokenizer(prompt, return_tensors="pt")["input_ids"][0]
            check_max_length(len(ids), max_length)
            new = self.group.generate({"model": model_name, "shard_ids": list(shard_ids), "ids": ids.tolist(),
                                       "max_length": int(max_length), "deadline": deadline, "gen": gen or {}})
            eng = self.loaded_models.get(f"{model_name}_{sorted(shard_ids)[0]}")
            new = apply_eos(torch.tensor([new], dtype=torch.int64), tuple(getattr(eng, "eos_token_ids", ())),
                            getattr(eng, "pad_token_id", None))
            return tokenizer.decode(torch.cat((ids.long(), new[0])), skip_special_tokens=True)

    return GroupWorker()


def main(argv=None):
    import argparse
    import threading

    from werkzeug.serving import make_server

    from .pipeline import init_distributed

    ap = argparse.ArgumentParser(description="N workers on N GPUs, shards joined by an NCCL pipeline; rank r "
                                             "serves the reference worker's routes on port + r")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=5000)
    a = ap.parse_args(argv)
    rank, world, local = init_distributed()
    ctrl = dist.new_group(backend="gloo")
    w = make_group_worker(rank, world, local, ctrl)
    srv = make_server(a.host, a.port + rank, w, threaded=True)
    if rank == 0:
        try:
            srv.serve_forever()
        finally:
            w.group.shutdown()
    else:
        threading.Thread(target=srv.serve_forever, daemon=True).start()
        torch.cuda.set_device(local)
        w.group.serve_forever()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
