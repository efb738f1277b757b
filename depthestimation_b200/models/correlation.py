"""Drop-ins for the reference's two correlation layers (forward): models/correlation_package/correlation.py:47-61
(FlowNet-style, any pad / kernel / displacement / strides) and its torch twin models/correlation_native.py:6-23
(kernel 1, strides 1, pad = max_displacement).  Both run the sm_100a kernel behind dpv_correlation.
Inference only: the reference's backward kernels (training) are outside the hot path."""
import torch

from .. import ops


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
        if input1.requires_grad or input2.requires_grad:
            raise NotImplementedError("Correlation: forward only (no autograd) in the B200 path")
        return ops.correlation(input1, input2, self.pad_size, self.kernel_size, self.max_displacement, self.stride1,
                               self.stride2)


class CorrelationNative(torch.nn.Module):
    """models/correlation_native.py: Correlation(max_displacement=4) -> [B,(2d+1)^2,H,W], mean over channels."""

    def __init__(self, max_displacement=4, *args, **kwargs):
        super().__init__()
        self.max_displacement = max_displacement
        self.output_dim = 2 * self.max_displacement + 1
        self.pad_size = self.max_displacement

    def forward(self, x1, x2):
        if x1.requires_grad or x2.requires_grad:
            raise NotImplementedError("Correlation: forward only (no autograd) in the B200 path")
        return ops.correlation(x1, x2, self.pad_size, 1, self.max_displacement, 1, 1)
