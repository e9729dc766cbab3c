"""neuralradiancecaching_b200 -- B200-native (sm_100a) neural-radiance-cache hot path behind a C-ABI.

csrc/      CUDA kernels + the C-ABI (include/nrc_b200.h)  -> libnrc_b200.so
api.py     NeuralRadianceCache, MRHE and the stand-alone operators
objects.py Camera, DirLight, PointLight, VolumeData, HdrEnvMap
renderer.py NrcHpmRenderer
loaders.py ReadFileDensity3D, ReadFileHdr4f (src/read_file.cpp); dist.py data-parallel plumbing
"""
from . import _lib
from ._lib import NrcError, PARAM_COUNT
from .api import (Context, NeuralRadianceCache, MRHE, mrhe_resolutions, encode, encode_dir, infer, train_accumulate,
                  optimizer_step, feature_tile_bytes, encode_tiles, mlp_forward_tiles)
from .objects import Camera, DirLight, PointLight, VolumeData, HdrEnvMap, hdr4f_to_cdf, vec_from_angles
from .renderer import NrcHpmRenderer
from .loaders import ReadFileDensity3D, ReadFileHdr4f
from .checkpoint import save_checkpoint, load_checkpoint
