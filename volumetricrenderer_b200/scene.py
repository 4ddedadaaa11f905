"""Host-side scene of the reference's main(): the two uniform blocks it writes every frame.

Mirrors /root/reference/source/Etna/TestMain.cpp:27-39 (struct layouts) and :222-245 (values):
    ObjectShaderData { mat4 Model, View, Projection }                  192 bytes
    GlobalShaderData { mat4 WorldToLocal; vec3 CameraPosition; mat4 MediaScroll }
                      140 bytes as the C++ compiler lays it out (NOT std140 — MediaScroll starts at
                      byte 76 while frag.glsl:9-14 reads it at byte 80)
The glm calls are restated from their closed forms in binary32 (glm is a system include of the
reference, not vendored): rotate, lookAtRH, perspectiveRH_ZO (VulkanHeader.h:9-10), inverse.
All matrices are column-major flat arrays m[col*4+row], like glm / GLSL.
"""
import numpy as np

f32 = np.float32


def _normalize(v):
    v = np.asarray(v, f32)
    inv = f32(1.0) / np.sqrt(f32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), dtype=f32)
    return (v * inv).astype(f32)


def _dot(a, b):
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def _cross(a, b):
    return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], f32)


def identity():
    return np.eye(4, dtype=f32).ravel()


def radians(deg):
    return f32(f32(deg) * f32(0.01745329251994329576923690768489))


def rotate(m, angle, axis):
    """glm::rotate(m, angle, axis) = m * R(angle, axis)   (TestMain.cpp:222,224)"""
    m = np.asarray(m, f32)
    c, s = f32(np.cos(f32(angle), dtype=f32)), f32(np.sin(f32(angle), dtype=f32))
    a = _normalize(axis)
    t = ((f32(1.0) - c) * a).astype(f32)
    R = np.zeros((3, 3), f32)
    R[0, 0] = c + t[0] * a[0]
    R[0, 1] = t[0] * a[1] + s * a[2]
    R[0, 2] = t[0] * a[2] - s * a[1]
    R[1, 0] = t[1] * a[0] - s * a[2]
    R[1, 1] = c + t[1] * a[1]
    R[1, 2] = t[1] * a[2] + s * a[0]
    R[2, 0] = t[2] * a[0] + s * a[1]
    R[2, 1] = t[2] * a[1] - s * a[0]
    R[2, 2] = c + t[2] * a[2]
    out = np.zeros(16, f32)
    for col in range(3):
        for r in range(4):
            out[col * 4 + r] = f32(f32(m[0 + r] * R[col, 0]) + f32(m[4 + r] * R[col, 1])) + f32(m[8 + r] * R[col, 2])
    out[12:16] = m[12:16]
    return out


def look_at(eye, center, up):
    """glm::lookAtRH (TestMain.cpp:225)"""
    eye, center, up = (np.asarray(v, f32) for v in (eye, center, up))
    f = _normalize(center - eye)
    s = _normalize(_cross(f, up))
    u = _cross(s, f)
    m = identity()
    m[0], m[4], m[8] = s
    m[1], m[5], m[9] = u
    m[2], m[6], m[10] = -f
    m[12] = -(s[0] * eye[0] + s[1] * eye[1] + s[2] * eye[2])
    m[13] = -(u[0] * eye[0] + u[1] * eye[1] + u[2] * eye[2])
    m[14] = f[0] * eye[0] + f[1] * eye[1] + f[2] * eye[2]
    return m


def perspective(fovy, aspect, z_near, z_far):
    """glm::perspectiveRH_ZO — GLM_FORCE_DEPTH_ZERO_TO_ONE (VulkanHeader.h:10; TestMain.cpp:226)"""
    t = f32(np.tan(f32(fovy) / f32(2.0), dtype=f32))
    zn, zf, aspect = f32(z_near), f32(z_far), f32(aspect)
    m = np.zeros(16, f32)
    m[0] = f32(1.0) / f32(aspect * t)
    m[5] = f32(1.0) / t
    m[10] = zf / f32(zn - zf)
    m[11] = f32(-1.0)
    m[14] = -f32(zf * zn) / f32(zf - zn)
    return m


def inverse(m):
    """general 4x4 inverse, formed in double and rounded once to binary32"""
    a = np.asarray(m, np.float64).reshape(4, 4).T          # column-major flat -> math matrix
    return np.linalg.inv(a).T.astype(f32).ravel()


def default_scene(phi_deg=0.0, theta_deg=0.0, frame_time=0.0, aspect=(1280, 720)):
    """-> (ubo0: float32[48] = 192 B, ubo1: float32[35] = the 140-byte C++ GlobalShaderData image).

    TestMain.cpp:222-245: Model = Rz(phi) * Ry(theta); View = lookAt((3,3,3), 0, +Z);
    Projection = perspective(45 deg, 1280/720, 0.1, 10) with [1][1] negated; WorldToLocal =
    inverse(Model); CameraPosition = (3,3,3); MediaScroll[0][0] = -frameTime (which the C++ layout
    puts at byte 76, i.e. into the std140 padding the shader never reads).
    """
    ubo0 = np.zeros(48, f32)
    rot = rotate(identity(), radians(phi_deg), (0.0, 0.0, 1.0))
    ubo0[0:16] = rotate(rot, radians(theta_deg), (0.0, 1.0, 0.0))
    ubo0[16:32] = look_at((3.0, 3.0, 3.0), (0.0, 0.0, 0.0), (0.0, 0.0, 1.0))
    proj = perspective(radians(45.0), f32(aspect[0]) / f32(aspect[1]), 0.1, 10.0)
    proj[5] *= f32(-1.0)
    ubo0[32:48] = proj
    ubo1 = np.zeros(35, f32)
    ubo1[0:16] = inverse(ubo0[0:16])
    ubo1[16:19] = 3.0
    ubo1[19] = -f32(frame_time)
    return ubo0, ubo1


def orbit_views(count):
    """The C5 view batch: the reference's A/D/W/S orbit (TestMain.cpp:177-184)
    sampled at phi = 5.625*k deg, theta = 20*sin(2*pi*k/count) deg.
    -> uint8[count, 336]: per view {ubo0[192], ubo1 as std140[144]} — the layout vr_params.view_ubos takes.
    """
    out = np.zeros((count, 336), np.uint8)
    for k in range(count):
        u0, u1 = default_scene(5.625 * k, 20.0 * np.sin(2.0 * np.pi * k / count))
        out[k, :192] = u0.view(np.uint8)
        out[k, 192:192 + 140] = u1.view(np.uint8)       # verbatim host bytes; the tail reads as zero
    return out
