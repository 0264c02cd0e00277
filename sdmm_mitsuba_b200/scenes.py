"""Analytic test scenes for the device path tracer (include/sdmm_b200_render.h): quads + a pinhole camera.

cornell_box() restates /root/reference/test-suite/scenes/cornell-box/cornell-box.xml: five `rectangle`
walls, two `cube`s and a `rectangle` area light, each with its toWorld matrix (:88-152), the `diffuse`
reflectances (:36-73), the 35 degree perspective sensor and its toWorld (:15-19), film 640 x 360 (:27-28).
Mitsuba's `rectangle` spans [-1,1]^2 in its local xy plane with normal +z, `cube` spans [-1,1]^3.
The light's radiance is given spectrally (400:0, 500:1600, 600:3180, 700:3680, :141); Mitsuba's
spectrum -> RGB conversion is not restated, the RGB triple below (200 x the classic cbox light
(18.387, 13.9873, 6.75357)) is FROZEN for this repository and used by the oracle renderer too."""
import numpy as np

LIGHT_RGB = (3677.4, 2797.46, 1350.714)


def _mat(s):
    return np.array([float(v) for v in s.split()], np.float64).reshape(4, 4)


def _quad(p0, e1, e2, facing, refl, emis=(0, 0, 0)):
    if np.dot(np.cross(e1, e2), facing) < 0:
        e1, e2 = e2, e1
    return dict(p0=np.asarray(p0, np.float32), e1=np.asarray(e1, np.float32), e2=np.asarray(e2, np.float32),
                reflectance=np.asarray(refl, np.float32), emission=np.asarray(emis, np.float32))


def rectangle(to_world, refl, flip=False, emis=(0, 0, 0)):
    m = _mat(to_world) if isinstance(to_world, str) else np.asarray(to_world, np.float64)
    c0, c1, o = m[:3, 0], m[:3, 1], m[:3, 3]
    facing = np.cross(c0, c1)            # Mitsuba: n = normalize(dpdu x dpdv), negated by flipNormals
    if flip:
        facing = -facing
    return [_quad(o - c0 - c1, 2 * c0, 2 * c1, facing, refl, emis)]


def cube(to_world, refl):
    m = _mat(to_world) if isinstance(to_world, str) else np.asarray(to_world, np.float64)
    cols, o = [m[:3, 0], m[:3, 1], m[:3, 2]], m[:3, 3]
    quads = []
    for a in range(3):
        b, c = (a + 1) % 3, (a + 2) % 3
        for sgn in (1.0, -1.0):
            centre = o + sgn * cols[a]
            quads.append(_quad(centre - cols[b] - cols[c], 2 * cols[b], 2 * cols[c], sgn * cols[a], refl))
    return quads


def cornell_box(width=640, height=360):
    white = (0.725, 0.71, 0.68)
    q = []
    q += rectangle("-4.37114e-008 1 4.37114e-008 0 0 -8.74228e-008 2 0 1 4.37114e-008 1.91069e-015 0 0 0 0 1", white)
    q += rectangle("-1 7.64274e-015 -1.74846e-007 0 8.74228e-008 8.74228e-008 -2 2 0 -1 -4.37114e-008 0 0 0 0 1", white)
    q += rectangle("1.91069e-015 1 1.31134e-007 0 1 3.82137e-015 -8.74228e-008 1 -4.37114e-008 1.31134e-007 -2 -1 0 0 0 1",
                   white, flip=True)
    q += rectangle("4.37114e-008 -1.74846e-007 2 1 1 3.82137e-015 -8.74228e-008 1 3.82137e-015 1 2.18557e-007 0 0 0 0 1",
                   (0.14, 0.45, 0.091), flip=True)
    q += rectangle("-4.37114e-008 8.74228e-008 -2 -1 1 3.82137e-015 -8.74228e-008 1 0 -1 -4.37114e-008 0 0 0 0 1",
                   (0.63, 0.065, 0.05), flip=True)
    q += cube("0.0851643 0.289542 1.31134e-008 0.328631 3.72265e-009 1.26563e-008 -0.3 0.3 -0.284951 0.0865363 "
              "5.73206e-016 0.374592 0 0 0 1", white)
    q += cube("0.286776 0.098229 -2.29282e-015 -0.335439 -4.36233e-009 1.23382e-008 -0.6 0.6 -0.0997984 0.282266 "
              "2.62268e-008 -0.291415 0 0 0 1", white)
    q += rectangle("4.700000e-02 -3.322060e-09 -1.561370e-09 -5.000000e-03 -4.108880e-09 7.806860e-10 -1.786000e-02 "
                   "1.980000e+00 4.108880e-09 3.800000e-02 1.661032e-09 -3.000000e-02 0 0 0 1", (0, 0, 0), emis=LIGHT_RGB)
    cam_m = _mat("-1 0 0 0 0 1 0 1 0 0 -1 6.8 0 0 0 1")
    camera = dict(origin=cam_m[:3, 3].astype(np.float32), right=cam_m[:3, 0].astype(np.float32),
                  up=cam_m[:3, 1].astype(np.float32), forward=cam_m[:3, 2].astype(np.float32), fov_x_degrees=35.0,
                  width=width, height=height)
    return q, camera


def scene_bounds(quads):
    """Scene AABB (getAABBWithoutCamera) and the largest extent = spatialNormalization
    (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:109-114)."""
    pts = []
    for q in quads:
        for u in (0, 1):
            for v in (0, 1):
                pts.append(q["p0"].astype(np.float64) + u * q["e1"] + v * q["e2"])
    pts = np.array(pts)
    lo, hi = pts.min(0), pts.max(0)
    return lo.astype(np.float32), hi.astype(np.float32), float((hi - lo).max())


def root_aabb(lo, hi, norm):
    """getAABB (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:314-332): normalised box padded by 1e-5."""
    eps = np.float32(1e-5)
    return (np.zeros(3, np.float32) - eps), ((hi - lo) / np.float32(norm)).astype(np.float32) + eps


def _box(lo, hi, refl):
    """Axis-aligned box [lo, hi] as six outward-facing quads."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    quads = []
    for a in range(3):
        b, c = (a + 1) % 3, (a + 2) % 3
        eb, ec = np.zeros(3), np.zeros(3)
        eb[b], ec[c] = hi[b] - lo[b], hi[c] - lo[c]
        for side, sgn in ((lo, -1.0), (hi, 1.0)):
            p0 = lo.copy()
            p0[a] = side[a]
            n = np.zeros(3)
            n[a] = sgn
            quads.append(_quad(p0, eb, ec, n, refl))
    return quads


def interior(width=640, height=360):
    """A procedurally generated three-room interior of analytic boxes (BASELINE.json configs[3] asks for a complex
    test-suite interior; every mesh under test-suite/scenes is a git-LFS pointer, so SURVEY.md section 8 F1 allows
    this instead).  Rooms of 2 x 2 x 2 in a row along x, joined by door openings in 0.1-thick dividing walls; the
    camera stands in the first room and looks down the row, the only light is a small ceiling panel in the LAST
    room, so the first room is lit by light that passed two doorways -- the case path guiding is for.
    61 quads: shell 6, dividing walls 2 x 3 boxes, three furniture boxes, one light."""
    white, warm, blue, green = (0.725, 0.71, 0.68), (0.7, 0.45, 0.25), (0.2, 0.3, 0.65), (0.14, 0.45, 0.091)
    q = []
    # shell [-3, 3] x [0, 2] x [-1, 1], normals facing inwards
    q.append(_quad((-3, 0, -1), (6, 0, 0), (0, 0, 2), (0, 1, 0), white))      # floor
    q.append(_quad((-3, 2, -1), (6, 0, 0), (0, 0, 2), (0, -1, 0), white))     # ceiling
    q.append(_quad((-3, 0, -1), (6, 0, 0), (0, 2, 0), (0, 0, 1), warm))       # back wall z = -1
    q.append(_quad((-3, 0, 1), (6, 0, 0), (0, 2, 0), (0, 0, -1), blue))       # front wall z = +1
    q.append(_quad((-3, 0, -1), (0, 0, 2), (0, 2, 0), (1, 0, 0), white))      # end wall x = -3
    q.append(_quad((3, 0, -1), (0, 0, 2), (0, 2, 0), (-1, 0, 0), green))      # end wall x = +3
    # dividing walls at x = -1 and x = +1 with a door (z in [-0.4, 0.4], y in [0, 1.5]); the second door is offset
    for xc, z0 in ((-1.0, -0.4), (1.0, 0.0)):
        z1 = z0 + 0.8
        q += _box((xc - 0.05, 0, -1), (xc + 0.05, 2, z0), white)
        q += _box((xc - 0.05, 0, z1), (xc + 0.05, 2, 1), white)
        q += _box((xc - 0.05, 1.5, z0), (xc + 0.05, 2, z1), white)
    # furniture
    q += _box((-2.2, 0, -0.8), (-1.6, 0.7, -0.3), white)
    q += _box((-0.5, 0, 0.3), (0.2, 1.1, 0.8), warm)
    q += _box((1.8, 0, -0.7), (2.5, 0.5, 0.0), white)
    # ceiling light in the last room, facing down
    q.append(_quad((1.9, 1.98, -0.25), (0.5, 0, 0), (0, 0, 0.5), (0, -1, 0), (0, 0, 0), emis=(400.0, 380.0, 330.0)))
    fwd, up = np.array([1.0, 0.0, 0.05]), np.array([0.0, 1.0, 0.0])
    fwd /= np.linalg.norm(fwd)
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    camera = dict(origin=np.array([-2.9, 1.0, 0.0], np.float32), right=right.astype(np.float32),
                  up=up.astype(np.float32), forward=fwd.astype(np.float32), fov_x_degrees=70.0, width=width, height=height)
    return q, camera
