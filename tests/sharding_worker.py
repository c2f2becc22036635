"""SURVEY.md §4 (iv) sharding determinism, run under `python -m torch.distributed.run --nproc-per-node N`:
every rank samples its contiguous shard of the prompt set (per-prompt x_T seeds), the uint8 images are all-gathered
over NCCL, and rank 0 re-samples EVERY other rank's shard on its own GPU with the same micro-batch size: the gathered
images must be bit-identical to the single-GPU ones (same kernels, same tile choices, fixed-order reductions)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from fgdm_b200 import _lib, synth
    from fgdm_b200 import golden_cases as gc
    from fgdm_b200.ldm import instantiate_from_config
    from fgdm_b200.pipeline import FGDMPipeline, all_gather_images, shard_indices
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    steps = int(os.environ.get("SHARD_STEPS", "12"))
    per = int(os.environ.get("SHARD_PER_RANK", "4"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    _lib.require_device()
    dist.init_process_group("nccl", device_id=dev)
    seg = instantiate_from_config(gc.fgdm_config())
    seg.device = dev
    seg.load_state_dict(synth.make_state_dict(seg.param_shapes(), gc.FGDM_WEIGHT_SEED))
    img = instantiate_from_config(gc.control_config())
    img.device = dev
    img.load_state_dict(synth.make_state_dict(img.param_shapes(), gc.CONTROL_WEIGHT_SEED))
    pipe = FGDMPipeline(seg, img, seg_steps=steps, img_steps=steps)

    def run(indices):
        d = {k: v.to(dev) for k, v in synth.prompt_inputs(indices, (64, 64)).items()}
        return pipe(d["c"], d["uc"], d["xT_seg"], d["c2"], d["uc2"], d["xT_img"])

    P = per * world
    mine = shard_indices(P, rank, world)
    gathered = all_gather_images(run(mine))
    assert gathered.shape[0] == P
    torch.cuda.synchronize()
    ok, detail = True, []
    if rank == 0:
        for r in range(world):
            idx = shard_indices(P, r, world)
            single = run(idx)
            eq = bool(torch.equal(single, gathered[idx[0]:idx[-1] + 1]))
            detail.append({"shard_of_rank": r, "prompts": idx, "bit_equal_to_single_gpu": eq,
                           "checksum": int(single.to(torch.int64).sum())})
            ok &= eq
        distinct = len({d["checksum"] for d in detail}) == world      # shards really differ from each other
        print(json.dumps({"sharding_determinism": ok and distinct, "world": world, "prompts": P, "ddim_steps": steps,
                          "whole_gather_checksum": int(gathered.to(torch.int64).sum()), "shards": detail}), flush=True)
        ok = ok and distinct
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
