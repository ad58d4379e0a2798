"""Deterministic synthetic soccar collision meshes in the reference's .cmf layout.

The real Rocket League arena meshes are not redistributable and are absent from the reference repo
(RocketSim::Init loads them from an external folder: RocketSim/src/RocketSim.cpp:70-100).  RocketSim's
soccar arena = 4 analytic planes (floor, ceiling, +-x walls: Sim/Arena/Arena.cpp:1060-1101) + triangle
meshes for everything else.  This module generates a stand-in for "everything else": back walls with goal
mouths, goal boxes, 45-degree corner walls and floor/wall fillet ramps.  The same bytes are fed to the
reference (RocketSim::InitFromMem, RocketSim.cpp:102-175) and to the CUDA library.

.cmf layout (RocketSim/src/CollisionMeshFile/CollisionMeshFile.cpp:11-35):
    int32 numTris, int32 numVerts, int32 tri[numTris][3], float vert[numVerts][3]   (Bullet units = uu/50)
"""
import numpy as np

UU_TO_BT = np.float32(1.0 / 50.0)

EXT_X, EXT_Y, HEIGHT = 4096.0, 5120.0, 2044.0
GOAL_HALF_W, GOAL_H, GOAL_DEPTH = 893.0, 642.775, 880.0
CORNER = 1152.0
RAMP_R = 256.0


class _Mesh:
    def __init__(self):
        self.verts = []
        self.tris = []

    def quad(self, p00, p10, p11, p01, nu=1, nv=1):
        """Grid of nu x nv cells over the bilinear patch p00-p10-p11-p01 (counter-clockwise seen from the
        side the normal points to)."""
        p00, p10, p11, p01 = [np.asarray(p, dtype=np.float64) for p in (p00, p10, p11, p01)]
        base = len(self.verts)
        for j in range(nv + 1):
            v = j / nv
            for i in range(nu + 1):
                u = i / nu
                p = (1 - u) * (1 - v) * p00 + u * (1 - v) * p10 + u * v * p11 + (1 - u) * v * p01
                self.verts.append(p)
        for j in range(nv):
            for i in range(nu):
                a = base + j * (nu + 1) + i
                b = a + 1
                c = a + (nu + 1) + 1
                d = a + (nu + 1)
                self.tris.append((a, b, c))
                self.tris.append((a, c, d))

    def blob(self):
        v = (np.asarray(self.verts, dtype=np.float64)).astype(np.float32) * UU_TO_BT
        t = np.asarray(self.tris, dtype=np.int32)
        hdr = np.asarray([len(t), len(v)], dtype=np.int32)
        return hdr.tobytes() + t.tobytes() + v.astype(np.float32).tobytes()


def _cells(length, target=768.0):
    return max(1, int(round(length / target)))


def _ramp(m, a, b, inward, segs=4):
    """Quarter-circle fillet of radius RAMP_R along the floor edge a->b (z=0 points on the wall line);
    `inward` is the unit horizontal vector pointing into the arena. Normals point into the arena/up."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    inward = np.asarray(inward, dtype=np.float64)
    n = _cells(np.linalg.norm(b - a))
    prev = None
    for s in range(segs + 1):
        ang = (np.pi / 2) * s / segs
        # centre of the fillet circle is RAMP_R inward and RAMP_R up
        off = inward * (RAMP_R - RAMP_R * np.sin(ang)) + np.array([0, 0, RAMP_R - RAMP_R * np.cos(ang)])
        cur = (a + off, b + off)
        if prev is not None:
            m.quad(prev[0], prev[1], cur[1], cur[0], nu=n, nv=1)
        prev = cur


def _orient(m, inside_pt=(0.0, 0.0, 1000.0)):
    """Flip triangles whose normal points away from the arena interior (consistent winding matters for
    Bullet's internal-edge classification)."""
    v = np.asarray(m.verts)
    inside = np.asarray(inside_pt)
    out = []
    for (a, b, c) in m.tris:
        n = np.cross(v[b] - v[a], v[c] - v[a])
        if np.dot(n, inside - v[a]) < 0:
            out.append((a, c, b))
        else:
            out.append((a, b, c))
    m.tris = out


def soccar_meshes():
    """Returns a list of .cmf blobs (bytes). One mesh per end (+y, -y) for walls and one per goal box,
    plus one for the side-wall ramps: 5 meshes."""
    blobs = []
    for sy in (+1.0, -1.0):
        m = _Mesh()
        Y = sy * EXT_Y
        xc = EXT_X - CORNER
        # back wall: left of goal, right of goal, above goal
        for (x0, x1, z0, z1) in ((-xc, -GOAL_HALF_W, 0.0, HEIGHT), (GOAL_HALF_W, xc, 0.0, HEIGHT),
                                 (-GOAL_HALF_W, GOAL_HALF_W, GOAL_H, HEIGHT)):
            m.quad((x0, Y, z0), (x1, Y, z0), (x1, Y, z1), (x0, Y, z1), nu=_cells(x1 - x0), nv=_cells(z1 - z0))
        # corner walls
        for sx in (+1.0, -1.0):
            p0 = (sx * xc, Y, 0.0)
            p1 = (sx * EXT_X, sy * (EXT_Y - CORNER), 0.0)
            L = np.hypot(CORNER, CORNER)
            m.quad(p0, p1, (p1[0], p1[1], HEIGHT), (p0[0], p0[1], HEIGHT), nu=_cells(L), nv=_cells(HEIGHT))
            inward = np.array([-sx, -sy, 0.0]) / np.sqrt(2.0)
            _ramp(m, p0, p1, inward)
        # back-wall ramps either side of the goal mouth
        _ramp(m, (-xc, Y, 0.0), (-GOAL_HALF_W, Y, 0.0), (0.0, -sy, 0.0))
        _ramp(m, (GOAL_HALF_W, Y, 0.0), (xc, Y, 0.0), (0.0, -sy, 0.0))
        _orient(m)
        blobs.append(m.blob())

        g = _Mesh()
        Yb = sy * (EXT_Y + GOAL_DEPTH)
        # goal box: two side walls, back wall, roof
        for sx in (+1.0, -1.0):
            g.quad((sx * GOAL_HALF_W, Y, 0.0), (sx * GOAL_HALF_W, Yb, 0.0), (sx * GOAL_HALF_W, Yb, GOAL_H),
                   (sx * GOAL_HALF_W, Y, GOAL_H), nu=2, nv=1)
        g.quad((-GOAL_HALF_W, Yb, 0.0), (GOAL_HALF_W, Yb, 0.0), (GOAL_HALF_W, Yb, GOAL_H), (-GOAL_HALF_W, Yb, GOAL_H), nu=2, nv=1)
        g.quad((-GOAL_HALF_W, Y, GOAL_H), (GOAL_HALF_W, Y, GOAL_H), (GOAL_HALF_W, Yb, GOAL_H), (-GOAL_HALF_W, Yb, GOAL_H), nu=2, nv=2)
        _orient(g, inside_pt=(0.0, sy * (EXT_Y + GOAL_DEPTH / 2), GOAL_H / 2))
        blobs.append(g.blob())

    s = _Mesh()
    yc = EXT_Y - CORNER
    for sx in (+1.0, -1.0):
        _ramp(s, (sx * EXT_X, -yc, 0.0), (sx * EXT_X, yc, 0.0), (-sx, 0.0, 0.0))
    _orient(s)
    blobs.append(s.blob())
    return blobs


def parse_cmf(blob):
    hdr = np.frombuffer(blob[:8], dtype=np.int32)
    nt, nv = int(hdr[0]), int(hdr[1])
    tris = np.frombuffer(blob[8:8 + nt * 12], dtype=np.int32).reshape(nt, 3)
    verts = np.frombuffer(blob[8 + nt * 12:8 + nt * 12 + nv * 12], dtype=np.float32).reshape(nv, 3)
    return tris, verts


if __name__ == "__main__":
    bl = soccar_meshes()
    for i, b in enumerate(bl):
        t, v = parse_cmf(b)
        print(i, "tris", len(t), "verts", len(v))
