"""TEST INFRASTRUCTURE ONLY — the backward chain of oracle.restatement in float64.

Same control flow as restatement.backward (reference app/deepdream.py:468-472: DConvolution2D.down :102-111, DDense.down
:311-321 with the zero-bias transposed kernel of :289-297, DActivation.down :251-261, DPooling.down :141-150, DFlatten.down
:355-366) with every multiply-add carried out in float64 on the float32 operands.  It exists to tell apart the rounding
of an implementation under test from the rounding of the float32 oracle itself: the forward tape
(activations, switches) is taken from restatement.forward unchanged.
Only tests/ and profiles/ diagnostics import this module; the product never does."""
import numpy as np
import torch
import torch.nn.functional as F


def backward64(entries, i, seed):
    g = torch.from_numpy(np.asarray(seed, np.float64))
    for j in range(i, -1, -1):
        e = entries[j]
        if e.kind == "conv":
            Wd = np.ascontiguousarray(np.transpose(e.W, (0, 1, 3, 2))[::-1, ::-1, :, :])
            wt = torch.from_numpy(Wd).double().permute(3, 2, 0, 1).contiguous()
            g = F.conv2d(g.permute(0, 3, 1, 2), wt, padding=1).permute(0, 2, 3, 1)
            if e.act == "relu":
                g = F.relu(g)
        elif e.kind == "dense":
            g = g @ torch.from_numpy(np.ascontiguousarray(e.W.T)).double()
        elif e.kind == "activation":
            if e.act == "relu":
                g = F.relu(g)
            elif e.act == "softmax":
                g = torch.softmax(g, dim=-1)
        elif e.kind == "pool":
            g = g.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2) * torch.from_numpy(e.switch)
        elif e.kind == "flatten":
            g = g.reshape([g.shape[0]] + list(e.shape))
    return g.numpy().squeeze()
