"""Scene objects and the renderer: the Python face of en::Camera, en::DirLight, en::PointLight, en::VolumeData
(+ vk::Texture3D), en::HdrEnvMap and en::NrcHpmRenderer over the C-ABI."""
import ctypes as C
import math

import numpy as np

from . import _lib as L
from .api import _f32, _ptr


# ---- glm restated (column-major 4x4, float32) -- src/Camera.cpp:164-179, src/DirLight.cpp:5-14 --------------------
def _perspective(fovy, aspect, z_near, z_far):
    """glm::perspective, right-handed, clip z in [-1, 1] (no GLM_FORCE_* defines in the reference)."""
    t = np.float32(math.tan(np.float32(fovy) / np.float32(2.0)))
    m = np.zeros((4, 4), np.float32)   # m[col][row]
    m[0][0] = np.float32(1.0) / (np.float32(aspect) * t)
    m[1][1] = np.float32(1.0) / t
    m[2][2] = -(np.float32(z_far) + np.float32(z_near)) / (np.float32(z_far) - np.float32(z_near))
    m[2][3] = np.float32(-1.0)
    m[3][2] = -(np.float32(2.0) * np.float32(z_far) * np.float32(z_near)) / (np.float32(z_far) - np.float32(z_near))
    return m


def _normalize(v):
    v = np.asarray(v, np.float32)
    return v / np.float32(np.sqrt(np.float32(np.dot(v, v))))


def _look_at(eye, center, up):
    """glm::lookAt (right-handed)."""
    eye, center, up = (np.asarray(a, np.float32) for a in (eye, center, up))
    f = _normalize(center - eye)
    s = _normalize(np.cross(f, up))
    u = np.cross(s, f)
    m = np.eye(4, dtype=np.float32)
    m[0][0], m[1][0], m[2][0] = s
    m[0][1], m[1][1], m[2][1] = u
    m[0][2], m[1][2], m[2][2] = -f
    m[3][0] = -np.dot(s, eye)
    m[3][1] = -np.dot(u, eye)
    m[3][2] = np.dot(f, eye)
    return m


def _rotate(angle, axis):
    """glm::rotate(angle, axis) (gtx/transform): column-major rotation matrix."""
    a = np.float32(angle)
    c, s = np.float32(math.cos(a)), np.float32(math.sin(a))
    ax = _normalize(axis)
    t = (np.float32(1.0) - c) * ax
    m = np.eye(4, dtype=np.float32)
    m[0][0] = c + t[0] * ax[0]; m[0][1] = t[0] * ax[1] + s * ax[2]; m[0][2] = t[0] * ax[2] - s * ax[1]
    m[1][0] = t[1] * ax[0] - s * ax[2]; m[1][1] = c + t[1] * ax[1]; m[1][2] = t[1] * ax[2] + s * ax[0]
    m[2][0] = t[2] * ax[0] + s * ax[1]; m[2][1] = t[2] * ax[1] - s * ax[0]; m[2][2] = c + t[2] * ax[2]
    return m


def _matmul_cm(a, b):
    """a * b for column-major matrices stored as m[col][row]."""
    return (a.T.astype(np.float32) @ b.T.astype(np.float32)).T.astype(np.float32)


def _apply_cm(m, v):
    return (m.T.astype(np.float32) @ np.asarray(v, np.float32)).astype(np.float32)


class Camera:
    """en::Camera(pos, viewDir, up, aspectRatio, fov, nearPlane, farPlane) -- Camera.hpp:24."""

    def __init__(self, pos, view_dir, up, aspect_ratio, fov, near_plane, far_plane):
        self.m_Pos = np.asarray(pos, np.float32)
        self.m_ViewDir = np.asarray(view_dir, np.float32)
        self.m_Up = np.asarray(up, np.float32)
        self.m_AspectRatio, self.m_Fov, self.m_NearPlane, self.m_FarPlane = aspect_ratio, fov, near_plane, far_plane
        self.projView = self.invProjView = None
        self.UpdateUniformBuffer()

    def UpdateUniformBuffer(self):   # Camera.cpp:164-173
        proj = _perspective(self.m_Fov, self.m_AspectRatio, self.m_NearPlane, self.m_FarPlane)
        view = _look_at(self.m_Pos, self.m_Pos + self.m_ViewDir, self.m_Up)
        self.projView = _matmul_cm(proj, view)
        # glm::inverse works in float; a float64 inverse rounded to float32 is the same matrix to ~1 ulp and the
        # matrices cross the boundary as explicit inputs anyway (glm is unpinned)
        self.invProjView = np.linalg.inv(self.projView.T.astype(np.float64)).T.astype(np.float32)

    def RotateAroundOrigin(self, axis, angle):   # Camera.cpp:175-179
        p = _apply_cm(_rotate(angle, axis), np.append(self.m_Pos, np.float32(1.0)))
        self.m_Pos = p[:3].astype(np.float32)
        self.m_ViewDir = -_normalize(self.m_Pos)

    def SetAspectRatio(self, width, height=None):
        self.m_AspectRatio = float(width) if height is None else float(width) / float(height)

    def GetPos(self):
        return self.m_Pos

    def data(self):
        c = L.CameraData()
        c.projView[:] = self.projView.reshape(16).tolist()
        c.invProjView[:] = self.invProjView.reshape(16).tolist()
        c.pos[:] = self.m_Pos.tolist()
        return c


def vec_from_angles(zenith, azimuth):
    """VecFromAngles (src/DirLight.cpp:5-14): rotate(azimuth, Y) * rotate(zenith, X) * (0, 1, 0, 1)."""
    m = _matmul_cm(_rotate(azimuth, (0.0, 1.0, 0.0)), _rotate(zenith, (1.0, 0.0, 0.0)))
    return _apply_cm(m, (0.0, 1.0, 0.0, 1.0))[:3]


class DirLight:
    """en::DirLight(zenith, azimuth, color, strength) -- DirLight.hpp:26."""

    def __init__(self, zenith, azimuth, color, strength):
        self.zenith, self.azimuth, self.color, self.strength = zenith, azimuth, np.asarray(color, np.float32), strength
        self.dir = vec_from_angles(zenith, azimuth)

    def SetZenith(self, z):
        self.zenith = z
        self.dir = vec_from_angles(self.zenith, self.azimuth)

    def SetAzimuth(self, a):
        self.azimuth = a
        self.dir = vec_from_angles(self.zenith, self.azimuth)

    def data(self):
        d = L.DirLightData()
        d.color[:] = self.color.tolist()
        d.zenith, d.azimuth, d.strength = self.zenith, self.azimuth, self.strength
        d.dir[:] = np.asarray(self.dir, np.float32).tolist()
        return d


class PointLight:
    """en::PointLight(pos, color, strength) -- PointLight.hpp:16."""

    def __init__(self, pos, color, strength):
        self.pos, self.color, self.strength = np.asarray(pos, np.float32), np.asarray(color, np.float32), strength

    def data(self):
        d = L.PointLightData()
        d.pos[:] = self.pos.tolist()
        d.color[:] = self.color.tolist()
        d.strength = self.strength
        return d


class VolumeData:
    """vk::Texture3D(density3D, LINEAR, CLAMP_TO_BORDER, OPAQUE_BLACK) + en::VolumeData(&tex) -- main.cpp:123-129.
    density: float32 [x][y][z] (the layout ReadFileDensity3D returns)."""

    def __init__(self, ctx, density_xyz):
        d = np.ascontiguousarray(density_xyz, dtype=np.float32)
        nx, ny, nz = d.shape
        self.ctx, self.shape = ctx, (nx, ny, nz)
        self.h = C.c_void_p()
        L.check(L.lib().nrc_volume_create(ctx.h, _ptr(d), C.c_uint32(nx), C.c_uint32(ny), C.c_uint32(nz), C.byref(self.h)))
        # defaults of src/VolumeData.cpp:82-89
        self.random = np.zeros(4, np.float32)
        self.useNN, self.showNonNN, self.densityFactor, self.g, self.noNnSpp, self.withNnSpp = 0, 0, 0.4, 0.7, 1, 1
        self._lcg = 0x2545F491

    def Update(self, camera_changed=False, random4=None):
        """VolumeData::Update (VolumeData.cpp:105-109): a new `random` vec4 per frame.  The reference draws it from
        glm::linearRand (unseeded); here it is an explicit input or a seeded LCG."""
        if random4 is None:
            r = []
            for _ in range(4):
                self._lcg = (self._lcg * 1664525 + 1013904223) & 0xFFFFFFFF
                r.append((self._lcg >> 8) / float(1 << 24))
            random4 = r
        self.random = np.asarray(random4, np.float32)
        self.push()

    def push(self):
        p = L.VolumeParams()
        p.random[:] = self.random.tolist()
        p.useNN, p.showNonNN, p.densityFactor, p.g = self.useNN, self.showNonNN, self.densityFactor, self.g
        p.noNnSpp, p.withNnSpp = self.noNnSpp, self.withNnSpp
        L.check(L.lib().nrc_volume_set_params(self.h, C.byref(p)))

    def sample(self, pos):
        pos = _f32(pos)
        out = np.empty(pos.shape[0], np.float32)
        L.check(L.lib().nrc_volume_sample(self.h, _ptr(pos), C.c_size_t(pos.shape[0]), _ptr(out)))
        return out

    def Destroy(self):
        if self.h:
            L.lib().nrc_volume_destroy(self.h)
            self.h = C.c_void_p()


class HdrEnvMap:
    """en::HdrEnvMap(width, height, hdr4f, cdfX, cdfY) -- HdrEnvMap.hpp:15-20."""

    def __init__(self, ctx, width, height, hdr4f, cdf_x=None, cdf_y=None):
        rgba = _f32(hdr4f)
        cx = None if cdf_x is None else _f32(cdf_x)
        cy = None if cdf_y is None else _f32(cdf_y)
        self.ctx, self.width, self.height = ctx, width, height
        self.h = C.c_void_p()
        L.check(L.lib().nrc_envmap_create(ctx.h, C.c_uint32(width), C.c_uint32(height), _ptr(rgba), _ptr(cx), _ptr(cy), C.byref(self.h)))
        self.directStrength, self.hpmStrength = 1.0, 8.0   # HdrEnvMap.cpp:108

    def set_importance_sampling(self, on):
        """Opt-in (not the shipped estimator): 4 of the 8 env directions per vertex from the inverse CDFs (fwd.frag:715, 727-744)."""
        L.check(L.lib().nrc_envmap_set_importance_sampling(self.h, C.c_int(1 if on else 0)))

    def set_strengths(self, direct, hpm):
        self.directStrength, self.hpmStrength = direct, hpm
        L.check(L.lib().nrc_envmap_set_strengths(self.h, C.c_float(direct), C.c_float(hpm)))

    def Destroy(self):
        if self.h:
            L.lib().nrc_envmap_destroy(self.h)
            self.h = C.c_void_p()


def hdr4f_to_cdf(ctx, hdr4f, width, height):
    """en::Hdr4fToCdf (src/read_file.cpp:141-206) -> [invCdfX (h*w), invCdfY (h)]."""
    rgba = _f32(hdr4f)
    cx = np.empty((height, width), np.float32)
    cy = np.empty(height, np.float32)
    L.check(L.lib().nrc_hdr4f_to_cdf(ctx.h, _ptr(rgba), C.c_uint32(width), C.c_uint32(height), _ptr(cx), _ptr(cy)))
    return [cx, cy]
