"""The 208-byte push-constant block of the compute pass, host side.

Field layout and default values are the reference's (src/vk_engine.h:145-178: CameraInfo,
EnvironmentData, RayTracerData, PushConstants; cameraAngles default src/vk_engine.h:325); the
camera matrix is built the way run_compute() builds it (src/vk_engine.cpp:1631-1653:
rotY * rotX * rotZ from Euler degrees, GLM column constructors).
"""
import numpy as np

PC_DTYPE = np.dtype([
    ("cameraRotation", "<f4", (4, 4)),      # column-major: [col][row]
    ("pos", "<f4", 3), ("nearPlane", "<f4"), ("aspectRatio", "<f4"), ("fov", "<f4"), ("_pad0", "<u4", 2),
    ("horizonColor", "<f4", 4), ("zenithColor", "<f4", 4), ("groundColor", "<f4", 3), ("_pad1", "<u4"),
    ("lightDir", "<f4", 4),
    ("progressive", "u1"), ("_p0", "u1", 3), ("singleRender", "u1"), ("_p1", "u1", 3),
    ("debug", "<i4"), ("raysPerPixel", "<u4"), ("bounceLimit", "<u4"), ("sphereCount", "<u4"),
    ("objectCount", "<u4"), ("triangleCap", "<u4"), ("boxCap", "<u4"), ("sampleLimit", "<u4"),
    ("frameCount", "<u4"), ("_pad2", "<u4"),
])
assert PC_DTYPE.itemsize == 208
assert PC_DTYPE.fields["horizonColor"][1] == 96 and PC_DTYPE.fields["progressive"][1] == 160
assert PC_DTYPE.fields["frameCount"][1] == 200


def _mat3_mul(a, b):
    # GLM mat3 * mat3 on column arrays a[col][row]: r[j][i] = a[0][i]*b[j][0] + a[1][i]*b[j][1] + a[2][i]*b[j][2]
    r = np.zeros((3, 3), np.float32)
    for j in range(3):
        for i in range(3):
            r[j][i] = np.float32(np.float32(a[0][i] * b[j][0] + a[1][i] * b[j][1]) + a[2][i] * b[j][2])
    return r


class PushConstants:
    """Mutable view of the block; `bytes(pc)` / `pc.buffer` is what b200pt_render takes."""

    def __init__(self, width, height, sphere_count, object_count):
        self.rec = np.zeros((), PC_DTYPE)
        r = self.rec
        r["pos"] = (0.0, -0.5, -3.5)
        r["nearPlane"] = 0.1
        r["aspectRatio"] = np.float32(width) / np.float32(height)      # run_compute, :1631
        r["fov"] = 50.0
        r["horizonColor"] = (0.986, 1.0, 0.902, 1000.0)                # w = sun focus
        r["zenithColor"] = (0.265, 0.595, 0.887, 10.0)                 # w = sun intensity
        r["groundColor"] = (0.431, 0.431, 0.431)
        d = np.array([2.0, 0.8, -3.0], np.float32)
        d = d * np.float32(1.0 / np.sqrt(np.float32(d @ d)))
        r["lightDir"] = (d[0], d[1], d[2], 0.0)                        # w = environment on
        r["progressive"] = 0
        r["singleRender"] = 0
        r["debug"] = -1
        r["raysPerPixel"] = 1
        r["bounceLimit"] = 8
        r["sphereCount"] = sphere_count                                # :1655
        r["objectCount"] = object_count                                # :1656
        r["triangleCap"] = 50
        r["boxCap"] = 200
        r["sampleLimit"] = 10
        r["frameCount"] = 0
        self.set_camera_angles(4.0, 0.0, 0.0)

    def set_camera_angles(self, ax, ay, az):
        tx, ty, tz = (np.float32(np.radians(np.float32(a))) for a in (ax, ay, az))
        c, s = (lambda t: np.float32(np.cos(t))), (lambda t: np.float32(np.sin(t)))
        rot_x = np.array([[1, 0, 0], [0, c(tx), -s(tx)], [0, s(tx), c(tx)]], np.float32)
        rot_y = np.array([[c(ty), 0, s(ty)], [0, 1, 0], [-s(ty), 0, c(ty)]], np.float32)
        rot_z = np.array([[c(tz), -s(tz), 0], [s(tz), c(tz), 0], [0, 0, 1]], np.float32)
        m = np.zeros((4, 4), np.float32)
        m[:3, :3] = _mat3_mul(_mat3_mul(rot_y, rot_x), rot_z)
        m[3, 3] = 1.0
        self.rec["cameraRotation"] = m

    def __getitem__(self, k):
        return self.rec[k]

    def __setitem__(self, k, v):
        self.rec[k] = v

    @property
    def buffer(self):
        return np.frombuffer(self.rec.tobytes(), np.uint8).copy()

    def __bytes__(self):
        return self.rec.tobytes()

    @classmethod
    def from_bytes(cls, b):
        self = cls.__new__(cls)
        self.rec = np.frombuffer(bytes(b), PC_DTYPE)[0].copy()
        return self
