"""One rank of the tensor-parallel parity run (launched by tests/test_gpu_tp.py under torch.distributed.run):
every rank builds its shard of the model, the group decodes greedily, and rank 0 compares logits and tokens with a
single-GPU engine on the same file."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "llama-fastapi-k8s-gpu_b200")]


def main():
    import torch
    import torch.distributed as dist
    from llama_cpp import synth
    from llama_cpp.engine import Engine, SamplingSettings
    from llama_cpp.gguf_file import GGUFFile
    arch, ftype = sys.argv[1], sys.argv[2]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    path = f"/tmp/tp_{arch}_{ftype}.gguf"
    if rank == 0:
        synth.write_synthetic_model(path, arch, ftype, seed=21, output_std=0.1, with_vocab=False)
    ids = [Engine.tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    g = GGUFFile(path)
    V = g.tensors["token_embd.weight"].shape[1]
    prompt = [int(t) for t in np.random.default_rng(3).integers(0, V, 6)]      # < 8 tokens: the single-GPU engine also takes the mat-vec path
    greedy = SamplingSettings(temperature=0.0, repeat_penalty=1.0)

    def run(eng):
        eng.ensure_pages(0, 48)
        eng.seq_begin(0, 0, [], greedy)
        eng.prefill(0, prompt)
        logits, toks = [eng.logits(0)], [eng.sample(0)]
        for _ in range(6):
            toks.append(eng.decode([0])[0])
            logits.append(eng.logits(0))
        toks += eng.decode_n(0, 5)                 # CUDA-graph replay with the collectives inside
        logits.append(eng.logits(0))
        return toks, logits

    eng = Engine(g, n_ctx=64, max_seqs=1, device=local, tp=(rank, world, ids[0]))
    if rank == 0:
        print(f"tp-mode {eng.mode()}", flush=True)
    toks, logits = run(eng)
    eng.close()
    all_toks = [None] * world
    dist.all_gather_object(all_toks, toks)
    assert all(t == all_toks[0] for t in all_toks), f"ranks disagree on the tokens: {all_toks}"
    if rank == 0:
        ref = Engine(g, n_ctx=64, max_seqs=1, device=local)
        rtoks, rlogits = run(ref)
        ref.close()
        sigma = float(rlogits[0].std())
        err = max(float(np.abs(a - b).max()) for a, b in zip(logits, rlogits))
        # teacher forcing is implicit: compare step by step while the tokens agree
        n_same = next((i for i, (a, b) in enumerate(zip(toks, rtoks)) if a != b), len(toks))
        print(f"tp-parity {arch} {ftype} world {world}: max |logit diff| {err:.4f} ({err / sigma:.2e} of the logit std), "
              f"tokens equal for {n_same}/{len(toks)} steps", flush=True)
        # same arithmetic, different summation order across ranks (and f16 KV entries that round apart): 5e-3 of the logit std
        assert n_same == len(toks) or err <= 5e-3 * sigma, (toks, rtoks)
        assert np.abs(logits[0] - rlogits[0]).max() <= 5e-3 * sigma
        print("tp-parity-ok", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
