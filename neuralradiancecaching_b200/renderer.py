"""en::NrcHpmRenderer over the C-ABI (include/engine/graphics/renderer/NrcHpmRenderer.hpp:20-40)."""
import ctypes as C

import numpy as np

from . import _lib as L
from .api import _ptr


class NrcHpmRenderer:
    """NrcHpmRenderer(width, height, trainWidth, trainHeight, camera, volumeData, dirLight, pointLight, hdrEnvMap, nrc,
    mrhe).  Like the reference the renderer keeps references to the seven objects; their current state is pushed
    before every Render()."""

    def __init__(self, width, height, train_width, train_height, camera, volume_data, dir_light, point_light, hdr_env_map,
                 nrc, mrhe):
        self.ctx = nrc.ctx
        self.width, self.height, self.trainW, self.trainH = width, height, train_width, train_height
        self.camera, self.volume, self.dirLight, self.pointLight, self.env, self.nrc, self.mrhe = (
            camera, volume_data, dir_light, point_light, hdr_env_map, nrc, mrhe)
        self.h = C.c_void_p()
        L.check(L.lib().nrc_renderer_create(self.ctx.h, C.c_uint32(width), C.c_uint32(height), C.c_uint32(train_width),
                                            C.c_uint32(train_height), volume_data.h, hdr_env_map.h, nrc.h, mrhe.h,
                                            C.byref(self.h)))

    def _push(self):
        cam, dl, pl = self.camera.data(), self.dirLight.data(), self.pointLight.data()
        L.check(L.lib().nrc_renderer_set_camera(self.h, C.byref(cam)))
        L.check(L.lib().nrc_renderer_set_dir_light(self.h, C.byref(dl)))
        L.check(L.lib().nrc_renderer_set_point_light(self.h, C.byref(pl)))
        self.volume.push()
        self.env.set_strengths(self.env.directStrength, self.env.hpmStrength)

    def Render(self, queue=None):
        """One frame: train -> nrc-step -> mrhe-step -> render (NrcHpmRenderer.cpp:57-72, 571-730); asynchronous."""
        self._push()
        L.check(L.lib().nrc_render(self.h))

    def FrameTrain(self):
        self._push()
        L.check(L.lib().nrc_frame_train(self.h))

    def FrameRender(self):
        self._push()
        L.check(L.lib().nrc_frame_render(self.h))

    def SetPartition(self, y0, y1, i0, i1):
        L.check(L.lib().nrc_renderer_set_partition(self.h, C.c_uint32(y0), C.c_uint32(y1), C.c_uint32(i0), C.c_uint32(i1)))

    def SetPartitionStrided(self, y_first, y_stride, i_first, i_stride):
        L.check(L.lib().nrc_renderer_set_partition_strided(self.h, C.c_uint32(y_first), C.c_uint32(y_stride), C.c_uint32(i_first),
                                                           C.c_uint32(i_stride)))

    def SetTrainEnvLight(self, on):
        """Opt-in: env-map term in the training targets (train.comp:1118 has it commented out)."""
        L.check(L.lib().nrc_renderer_set_train_env_light(self.h, C.c_int(1 if on else 0)))

    def GetPoolCapacity(self):
        """Vertices the render pool of the last frame could hold (the library's choice unless SetPoolCapacity was called)."""
        v = C.c_uint64(0)
        L.check(L.lib().nrc_renderer_get_pool_capacity(self.h, C.byref(v)))
        return int(v.value)

    def SetPoolCapacity(self, vertices):
        """Initial render vertex pool capacity (0 = default); an overflowing frame is redone with the worst-case pool."""
        L.check(L.lib().nrc_renderer_set_pool_capacity(self.h, C.c_uint64(vertices)))

    def partition_bytes(self, fmt=L.FRAME_BGRA8):
        return int(L.lib().nrc_frame_partition_bytes(self.h, C.c_int(fmt)))

    def read_frame_partition(self, out, fmt=L.FRAME_BGRA8):
        """This rank's rows only, compacted into `out` (partition_bytes() bytes)."""
        L.check(L.lib().nrc_read_frame_partition(self.h, _ptr(out), C.c_int(fmt)))
        return out

    def ResizeFrame(self, width, height):
        L.check(L.lib().nrc_resize(self.h, C.c_uint32(width), C.c_uint32(height)))
        self.width, self.height = width, height

    def GetImageDataSize(self):
        return int(L.lib().nrc_frame_bytes(self.h, C.c_int(L.FRAME_BGRA8)))

    def GetImage(self, fmt=L.FRAME_RGBA32F):
        """Device address of the frame (stands in for the VkImage)."""
        return L.lib().nrc_frame_ptr(self.h, C.c_int(fmt))

    def read_frame_partition_async(self, out, fmt=L.FRAME_BGRA8):
        """Queue the copy of this rank's rows into `out` (page-locked host memory); returns a ticket for readback_wait()."""
        t = C.c_int(-1)
        L.check(L.lib().nrc_read_frame_partition_async(self.h, _ptr(out), C.c_int(fmt), C.byref(t)))
        return t.value

    def readback_wait(self, ticket):
        L.check(L.lib().nrc_readback_wait(self.h, C.c_int(ticket)))

    def read_frame(self, fmt=L.FRAME_RGBA32F, out=None):
        if out is None:
            out = (np.empty((self.height, self.width, 4), np.float32) if fmt == L.FRAME_RGBA32F
                   else np.empty((self.height, self.width, 4), np.uint8))
        L.check(L.lib().nrc_read_frame(self.h, _ptr(out), C.c_int(fmt)))
        return out

    def read_render_probes(self):
        cnt = np.empty((self.height, self.width, 4), np.uint32)
        qry = np.empty((self.height, self.width, 8), np.float32)
        L.check(L.lib().nrc_read_render_probes(self.h, _ptr(cnt), _ptr(qry)))
        return cnt, qry

    def read_train_probes(self):
        n = self.trainW * self.trainH
        rec = np.zeros((n, 9), np.float32)
        cnt = np.zeros((n, 4), np.uint32)
        L.check(L.lib().nrc_read_train_probes(self.h, _ptr(rec), _ptr(cnt)))
        return rec, cnt

    def stats(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        L.check(L.lib().nrc_frame_stats(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"infer_queries": a.value, "train_samples": b.value, "density_taps": c.value}

    def totals(self, reset=False):
        t = (C.c_uint64 * 6)()
        L.check(L.lib().nrc_frame_totals(self.h, t, C.c_int(1 if reset else 0)))
        keys = ("infer_queries", "train_samples", "render_taps", "train_taps", "render_vertices", "train_vertices")
        return dict(zip(keys, (int(x) for x in t)))

    def Destroy(self):
        if self.h:
            L.lib().nrc_renderer_destroy(self.h)
            self.h = C.c_void_p()
