"""Convert between a TF1-layout checkpoint (model.ckpt-N.index / .data-*) and a torch file holding
{variable name: tensor} (torch.save of a plain dict), in either direction.

  python -m distributed_tensorflow_framework_b200.cli.convert_checkpoint <model_dir | prefix> out.pt
  python -m distributed_tensorflow_framework_b200.cli.convert_checkpoint in.pt <new prefix, e.g. dir/model.ckpt-0>

Variable names, shapes and dtypes are kept (bfloat16 tensors are stored as bfloat16 on both sides); `global_step`
travels as an int64 scalar. The checkpoint direction also writes the `checkpoint` state file next to the prefix.
"""
import os
import sys

import numpy as np
import torch

from ..io import tensorbundle as tb


def checkpoint_to_torch(prefix, out_path):
    state = {}
    with tb.BundleReader(prefix) as r:
        for name in r.keys():
            v = r.get_tensor(name)
            if r.dtype_of(name) == tb.DT_BFLOAT16:  # raw 16-bit patterns
                state[name] = torch.from_numpy(np.ascontiguousarray(v).view(np.int16)).view(torch.bfloat16)
            else:
                state[name] = torch.from_numpy(np.array(v))
    torch.save(state, out_path)
    return state


def torch_to_checkpoint(in_path, prefix):
    state = torch.load(in_path, map_location="cpu")
    if not isinstance(state, dict):
        raise TypeError("%s does not hold a dict of tensors" % in_path)
    tensors = {}
    for name, t in state.items():
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(t)
        t = t.detach().cpu().contiguous()
        if t.dtype == torch.bfloat16:
            tensors[name] = (t.view(torch.int16).numpy().view(np.uint16), "bfloat16")
        else:
            tensors[name] = t.numpy()
    tb.write_bundle(prefix, tensors)
    model_dir = os.path.dirname(os.path.abspath(prefix))
    name = os.path.basename(prefix)
    _, all_paths = tb.read_checkpoint_state(model_dir)
    tb.write_checkpoint_state(model_dir, name, [p for p in all_paths if p != name] + [name])
    return sorted(tensors)


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) != 2:
        print(__doc__)
        return 2
    src, dst = argv
    if src.endswith((".pt", ".pth")):
        names = torch_to_checkpoint(src, dst)
        print("wrote %d tensors to %s" % (len(names), dst))
        return 0
    prefix = tb.latest_checkpoint(src) if os.path.isdir(src) else src
    if prefix is None:
        print("no checkpoint in %s" % src, file=sys.stderr)
        return 1
    state = checkpoint_to_torch(prefix, dst)
    print("wrote %d tensors from %s to %s" % (len(state), prefix, dst))
    return 0


if __name__ == "__main__":
    sys.exit(main())
