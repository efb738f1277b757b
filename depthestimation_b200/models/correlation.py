"""Drop-ins for the reference's two correlation layers: models/correlation_package/correlation.py:6-61
(FlowNet-style CorrelationFunction / Correlation, any pad / kernel / displacement / strides) and its torch twin
models/correlation_native.py:6-23 (kernel 1, strides 1, pad = max_displacement).  Forward and backward run the sm_100a
kernels behind dpv_correlation / dpv_correlation_backward; the backward is defined for stride1 = 1 like the
reference's (its launch writes out of bounds otherwise, correlation_cuda_kernel.cu:130-131,486)."""
import torch

from .. import ops


class CorrelationFunction(torch.autograd.Function):
    """correlation.py:6-45 as a new-style autograd Function: apply(input1, input2, pad_size, kernel_size,
    max_displacement, stride1, stride2)."""

    @staticmethod
    def forward(ctx, input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2):
        ctx.save_for_backward(input1, input2)
        ctx.cfg = (pad_size, kernel_size, max_displacement, stride1, stride2)
        return ops.correlation(input1, input2, pad_size, kernel_size, max_displacement, stride1, stride2)

    @staticmethod
    def backward(ctx, grad_output):
        input1, input2 = ctx.saved_tensors
        g1, g2 = ops.correlation_backward(grad_output.contiguous(), input1, input2, *ctx.cfg)
        return g1, g2, None, None, None, None, None


class Correlation(torch.nn.Module):
    def __init__(self, pad_size=0, kernel_size=0, max_displacement=0, stride1=1, stride2=2, corr_multiply=1):
        super().__init__()
        self.pad_size = pad_size
        self.kernel_size = kernel_size
        self.max_displacement = max_displacement
        self.stride1 = stride1
        self.stride2 = stride2
        self.corr_multiply = corr_multiply          # accepted and, as in the reference's kernels, not used

    def forward(self, input1, input2):
        return CorrelationFunction.apply(input1, input2, self.pad_size, self.kernel_size, self.max_displacement,
                                         self.stride1, self.stride2)


class CorrelationNative(torch.nn.Module):
    """models/correlation_native.py: Correlation(max_displacement=4) -> [B,(2d+1)^2,H,W], mean over channels."""

    def __init__(self, max_displacement=4, *args, **kwargs):
        super().__init__()
        self.max_displacement = max_displacement
        self.output_dim = 2 * self.max_displacement + 1
        self.pad_size = self.max_displacement

    def forward(self, x1, x2):
        return CorrelationFunction.apply(x1, x2, self.pad_size, 1, self.max_displacement, 1, 1)
