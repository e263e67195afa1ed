#!/usr/bin/env python
"""Under torchrun on >= 2 GPUs: does torch's symmetric memory give this box a multicast (NVLS) mapping?
Prints what a multimem all-reduce kernel would need: multicast pointer, peer buffer pointers, signal pads."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"rank": rank, "world": world}
    try:
        import torch.distributed._symmetric_memory as symm

        t = symm.empty(1 << 20, dtype=torch.bfloat16, device=dev)
        try:
            hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
        except TypeError:
            hdl = symm.rendezvous(t, group=dist.group.WORLD)
        out["multicast_ptr"] = int(getattr(hdl, "multicast_ptr", 0) or 0)
        out["buffer_ptrs"] = [int(p) for p in hdl.buffer_ptrs]
        out["signal_pad_ptrs"] = [int(p) for p in hdl.signal_pad_ptrs]
        out["signal_pad_size"] = int(hdl.signal_pad_size)
        out["buffer_size"] = int(hdl.buffer_size)
        try:
            out["has_multicast_support"] = bool(type(hdl).has_multicast_support(torch._C._autograd.DeviceType.CUDA, local))
        except Exception as e:  # noqa: BLE001
            out["has_multicast_support"] = repr(e)[:120]
        # smoke: torch's own multimem all-reduce on the symmetric tensor
        t.fill_(rank + 1)
        dist.barrier()
        try:
            torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
            torch.cuda.synchronize()
            out["torch_multimem_all_reduce"] = float(t[0].item())
        except Exception as e:  # noqa: BLE001
            out["torch_multimem_all_reduce"] = repr(e)[:200]
    except Exception as e:  # noqa: BLE001
        out["error"] = repr(e)[:400]
    print("NVLS_PROBE " + json.dumps(out), flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)


if __name__ == "__main__":
    main()
