"""B200-native multi-view deconvolution inner loop behind the reference's native hook.

Package layout: csrc/ (CUDA kernels + C ABI -> libmvdecon.so), _lib (ctypes binding),
cuda (mirror of spim.process.cuda), deconvolution (mirror of spim.process.fusion.deconvolution).
"""
from . import _lib
from ._lib import MVDeconError, load
from .cuda import (Block, BlockGeneratorFixedSizePrecise, CUDADevice, CUDAFourierConvolution, CUDASeparableConvolution,
                   CUDASeparableConvolutionFunctions, CUDATools, convolve, plan_tiles)
from .fusion import ExtractPSF, VirtualWeights, WeightNormalizer, transformInputAndWeights
from .deconvolution import AdjustInput, Mirror, MVDeconFFT, MVDeconInput, MVDeconvolution, PSFTYPE

__all__ = ["ExtractPSF", "VirtualWeights", "WeightNormalizer", "transformInputAndWeights", "MVDeconError", "load", "Block", "BlockGeneratorFixedSizePrecise", "CUDADevice", "CUDAFourierConvolution", "CUDASeparableConvolution", "CUDASeparableConvolutionFunctions", "CUDATools",
           "convolve", "plan_tiles", "AdjustInput", "Mirror", "MVDeconFFT", "MVDeconInput", "MVDeconvolution", "PSFTYPE"]
