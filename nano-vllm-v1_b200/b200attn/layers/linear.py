"""Drop-in for the one collective next to the attention path: RowParallelLinear.forward
(nanovllm/layers/linear.py:335-411).  Same arithmetic and bias rule (bias only on rank 0, so it is
added once); only the all-reduce of decode-sized messages changes carrier: peer-memory
kernel now, or -- inside a model forward -- fused into the add+RMSNorm launch that follows (NVLS multimem)."""
import torch.nn.functional as F

from .. import ops


def row_parallel_forward(self, x):
    y = F.linear(x, self.weight, self.bias if self.tp_rank == 0 else None)
    if self.tp_size > 1:
        y = ops.tp_row_parallel_output(y)          # all-reduced now, or by the RMSNorm launch that consumes it
    return y
