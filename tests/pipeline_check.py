"""torchrun entry: N pipeline stages over NCCL == one full-range engine, token for token.
Launched by tests/test_pipeline_gpu.py (needs >= 2 GPUs)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden_util import load_case, spec_kwargs  # noqa: E402

from dli_b200.engine import ModelSpec, ShardEngine  # noqa: E402
from dli_b200.pipeline import (DistExchange, EngineStage, MailboxRunner, MailStage, PipelineRunner,  # noqa: E402
                               VocabParallelRunner, build_stage_engine, init_distributed, mailbox_period,
                               run_chunked_prefill)
from oracle import decoder_ref as R  # noqa: E402


def main():
    rank, world, local = init_distributed()
    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    B, T = c["prompts"].shape
    rounds, M, b = 6, B, 1  # one sequence per micro-batch slot
    vp = os.environ.get("DLI_PIPE_VP", "0") == "1"  # vocabulary-parallel head
    eng = build_stage_engine(spec, rank, world, local, max_batch=M * b, max_seq_len=128, vocab_parallel=vp)
    eng.load_state_dict(sd)
    chunks = int(os.environ.get("DLI_CHECK_CHUNKS", "1"))  # > 1: chunked, pipelined prefill
    stage = EngineStage(eng, b, T, T + rounds + 2, M, prefill_chunks=chunks)
    if stage.is_first:
        for mb in range(M):
            stage.prompt_ids[mb] = c["prompts"][mb].to("cuda", torch.int32).contiguous()
    runner = (VocabParallelRunner if vp else PipelineRunner)(stage, rank, world, M)
    run_chunked_prefill(stage, rank, world, M, stage.chunks)
    runner.run("prefill")
    first = None
    if stage.is_last and not vp:
        first = torch.stack([eng._state(b, mb)["tokens"].clone() for mb in range(M)])  # [M, b]
    hop = int(os.environ.get("DLI_CHECK_MAIL", "0"))  # > 0: decode over per-tick mailboxes with this hop
    if hop:
        P = mailbox_period(M, world, hop, vp)
        ms = MailStage(eng, b, M, P, hop, rank, world)
        comm = torch.cuda.Stream() if hop > 1 else None  # hop 2: the all-gather runs under the next tick
        MailboxRunner(ms, rank, world, M, hop=hop, vp=vp, exchange=DistExchange(comm_stream=comm)).run(rounds)
    else:
        runner.run("decode", rounds)
    torch.cuda.synchronize()
    ok = True
    holder = stage.is_first if vp else stage.is_last  # who ends up with the tokens
    if holder:
        logs = torch.stack([eng._state(b, mb)["token_log"][:rounds, 0] for mb in range(M)])  # [M, rounds]
        if vp:  # stage 0 logged the token it consumed at each step; the last choice is still in `tokens`
            final = torch.stack([eng._state(b, mb)["tokens"].clone() for mb in range(M)])
            got = torch.cat((logs, final[:, :1]), dim=1).cpu().long()
        else:
            got = torch.cat((first[:, :1], logs), dim=1).cpu().long()  # [M, 1 + rounds]
        full = ShardEngine(spec, device=f"cuda:{local}", max_batch=B, max_seq_len=128)
        full.load_state_dict(sd)
        if stage.chunks > 1:
            # same chunking and the same rows per call on the single engine (one sequence at a time, like a
            # slot): the projections pick their kernel by row count, so only then are the bits identical
            rows = []
            for mb in range(M):
                full.start_batch(1, T, T + rounds + 2)
                toks = [full.prefill_chunked(1, T, stage.Tc, ids=c["prompts"][mb].to("cuda", torch.int32)).clone()]
                for _ in range(rounds):
                    toks.append(full.decode_step(1).clone())
                rows.append(torch.cat(toks))
            want = torch.stack(rows).cpu().long()
        else:
            want = full.generate(c["prompts"], T + 1 + rounds)[:, T:]
        ok = torch.equal(got, want)
        print("pipeline tokens", got.tolist(), "single-engine tokens", want.tolist(), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PIPELINE_OK" if flag.item() == 1 else "PIPELINE_MISMATCH", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
