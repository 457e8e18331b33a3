"""GPUDetector = GPUGeometry + channel map and DAQ CDFs (``chroma/gpu/detector.py:14-39``)."""
import numpy as np

from chroma_b200 import _lib
from chroma_b200.gpu.geometry import GPUGeometry
from chroma_b200.gpu.tools import to_gpu


class GPUDetector(GPUGeometry):
    def __init__(self, detector, wavelengths=None, print_usage=False, instanced=False):
        GPUGeometry.__init__(self, detector, wavelengths=wavelengths, print_usage=False, instanced=instanced)
        s2c = _lib.carray(detector.solid_id_to_channel_index, np.int32)
        tx = _lib.carray(detector.time_cdf[0], np.float32)
        ty = _lib.carray(detector.time_cdf[1], np.float32)
        qx = _lib.carray(detector.charge_cdf[0], np.float32)
        qy = _lib.carray(detector.charge_cdf[1], np.float32)
        if len(tx) != len(ty) or len(qx) != len(qy):
            raise ValueError('time/charge CDF x and y arrays must have the same length')
        self.nchannels = detector.num_channels()
        _lib.call('chroma_detector_attach', self.handle, _lib.ptr(s2c), len(s2c), self.nchannels,
                  _lib.ptr(tx), _lib.ptr(ty), len(tx), _lib.ptr(qx), _lib.ptr(qy), len(qx))
        # attributes of the reference class (host mirrors uploaded for callers that want them)
        self.solid_id_to_channel_index_gpu = to_gpu(s2c, ctx=self.ctx)
        self.time_cdf_x_gpu, self.time_cdf_y_gpu = to_gpu(tx, ctx=self.ctx), to_gpu(ty, ctx=self.ctx)
        self.charge_cdf_x_gpu, self.charge_cdf_y_gpu = to_gpu(qx, ctx=self.ctx), to_gpu(qy, ctx=self.ctx)
        self.detector_gpu = self.handle
