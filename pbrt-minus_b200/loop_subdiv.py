"""Shape "loopsubdiv": Loop subdivision surfaces as a host mesh producer.

Restates LoopSubdivide (shapes/loop_subdivide.cu:183-434; used by Shape::create, base/shape.cu:65-74): `levels` rounds of
Loop refinement (even vertices: one-ring rule with beta = 1/16 for regular vertices, 3/(8 valence) otherwise, 3/16 for
valence 3, boundary rule 1/8; odd vertices: 3/8 3/8 1/8 1/8, or the midpoint on a boundary edge), then every vertex is
pushed to the limit surface (gamma = 1 / (valence + 3 / (8 beta)); boundary 1/5) and gets the limit-surface normal S x T.
The result feeds the flat triangle scene like any "trianglemesh": P = limit positions, N = the (unnormalised) normals.

Arithmetic is fp32 in the reference's operation order, vertex and face numbering follow the reference (children of the
old vertices first, then the edge vertices in face / edge order; four children per face), so the mesh that reaches the BVH
build is the same mesh.  The mesh is held as index arrays (face -> 3 vertices, face -> 3 neighbours across edge k = (k, k+1),
vertex -> one incident face), not as a pointer graph.
"""
import math

import numpy as np

F32 = np.float32


def _nxt(i):
    return (i + 1) % 3


def _prv(i):
    return (i + 2) % 3


class _Mesh:
    """one refinement level"""

    def __init__(self, n_verts, n_faces):
        self.p = np.zeros((n_verts, 3), dtype=F32)
        self.start = [-1] * n_verts          # one face around the vertex
        self.regular = [False] * n_verts
        self.boundary = [False] * n_verts
        self.fv = [[-1, -1, -1] for _ in range(n_faces)]   # vertices of a face
        self.ff = [[-1, -1, -1] for _ in range(n_faces)]   # neighbour across the edge (k, k+1)

    def vnum(self, f, v):
        return self.fv[f].index(v)

    def next_face(self, f, v):
        return self.ff[f][self.vnum(f, v)]

    def prev_face(self, f, v):
        return self.ff[f][_prv(self.vnum(f, v))]

    def next_vert(self, f, v):
        return self.fv[f][_nxt(self.vnum(f, v))]

    def prev_vert(self, f, v):
        return self.fv[f][_prv(self.vnum(f, v))]

    def other_vert(self, f, a, b):
        for w in self.fv[f]:
            if w != a and w != b:
                return w
        return -1

    def valence(self, v):
        f = self.start[v]
        n = 1
        if not self.boundary[v]:
            while True:
                f = self.next_face(f, v)
                if f == self.start[v]:
                    return n
                n += 1
        while True:
            f = self.next_face(f, v)
            if f < 0:
                break
            n += 1
        f = self.start[v]
        while True:
            f = self.prev_face(f, v)
            if f < 0:
                break
            n += 1
        return n + 1

    def one_ring(self, v):
        ring = []
        if not self.boundary[v]:
            f = self.start[v]
            while True:
                ring.append(self.next_vert(f, v))
                f = self.next_face(f, v)
                if f == self.start[v]:
                    return ring
        f = self.start[v]
        while True:
            f2 = self.next_face(f, v)
            if f2 < 0:
                break
            f = f2
        ring.append(self.next_vert(f, v))
        while f >= 0:
            ring.append(self.prev_vert(f, v))
            f = self.prev_face(f, v)
        return ring

    def weight_one_ring(self, v, beta):
        ring = self.one_ring(v)
        valence = len(ring)
        beta = F32(beta)
        p = F32(F32(1) - F32(F32(valence) * beta)) * self.p[v]
        for w in ring:
            p = p + beta * self.p[w]
        return p.astype(F32)

    def weight_boundary(self, v, beta):
        ring = self.one_ring(v)
        beta = F32(beta)
        p = F32(F32(1) - F32(F32(2) * beta)) * self.p[v]
        p = p + beta * self.p[ring[0]]
        p = p + beta * self.p[ring[-1]]
        return p.astype(F32)


def _beta(valence):
    return F32(3) / F32(16) if valence == 3 else F32(3) / F32(F32(8) * F32(valence))


def _gamma(valence):
    return F32(1) / F32(F32(valence) + F32(F32(3) / F32(F32(8) * _beta(valence))))


def loop_subdivide(levels, indices, P):
    """-> (P_limit [n, 3] f32, indices [m, 3] i32, normals [n, 3] f32)"""
    P = np.asarray(P, dtype=F32).reshape(-1, 3)
    idx = np.asarray(indices, dtype=np.int64).reshape(-1, 3)
    m = _Mesh(len(P), len(idx))
    m.p[:] = P
    for f, tri in enumerate(idx):
        for j in range(3):
            m.fv[f][j] = int(tri[j])
            m.start[int(tri[j])] = f
    # neighbours: the second face to name an (unordered) edge closes it
    open_edges = {}
    for f in range(len(idx)):
        for k in range(3):
            a, b = m.fv[f][k], m.fv[f][_nxt(k)]
            key = (a, b) if a < b else (b, a)
            first = open_edges.pop(key, None)
            if first is None:
                open_edges[key] = (f, k)
            else:
                m.ff[first[0]][first[1]] = f
                m.ff[f][k] = first[0]
    for v in range(len(P)):
        if m.start[v] < 0:
            raise ValueError("loopsubdiv: vertex %d belongs to no face" % v)
        f = m.start[v]
        while True:
            f = m.next_face(f, v)
            if f < 0 or f == m.start[v]:
                break
        m.boundary[v] = f < 0
        val = m.valence(v)
        m.regular[v] = (val == 6) if not m.boundary[v] else (val == 4)

    for _ in range(max(int(levels), 0)):
        nv, nf = len(m.p), len(m.fv)
        # odd vertices are numbered in face / edge order after the nv even ones
        edge_vert = {}
        odd_p, odd_boundary, odd_start = [], [], []
        for f in range(nf):
            for k in range(3):
                a, b = m.fv[f][k], m.fv[f][_nxt(k)]
                key = (a, b) if a < b else (b, a)
                if key in edge_vert:
                    continue
                edge_vert[key] = nv + len(odd_p)
                on_boundary = m.ff[f][k] < 0
                if on_boundary:
                    p = F32(0.5) * m.p[a]
                    p = p + F32(0.5) * m.p[b]
                else:
                    p = F32(0.375) * m.p[a]
                    p = p + F32(0.375) * m.p[b]
                    p = p + F32(0.125) * m.p[m.other_vert(f, a, b)]
                    p = p + F32(0.125) * m.p[m.other_vert(m.ff[f][k], a, b)]
                odd_p.append(p.astype(F32))
                odd_boundary.append(on_boundary)
                odd_start.append(4 * f + 3)
        c = _Mesh(nv + len(odd_p), 4 * nf)
        for v in range(nv):
            c.regular[v] = m.regular[v]
            c.boundary[v] = m.boundary[v]
            if not m.boundary[v]:
                c.p[v] = m.weight_one_ring(v, F32(1) / F32(16) if m.regular[v] else _beta(m.valence(v)))
            else:
                c.p[v] = m.weight_boundary(v, F32(1) / F32(8))
            c.start[v] = 4 * m.start[v] + m.vnum(m.start[v], v)
        for i, p in enumerate(odd_p):
            c.p[nv + i] = p
            c.regular[nv + i] = True
            c.boundary[nv + i] = odd_boundary[i]
            c.start[nv + i] = odd_start[i]
        for f in range(nf):
            for j in range(3):
                # siblings
                c.ff[4 * f + 3][j] = 4 * f + _nxt(j)
                c.ff[4 * f + j][_nxt(j)] = 4 * f + 3
                # children of the neighbours
                f2 = m.ff[f][j]
                c.ff[4 * f + j][j] = 4 * f2 + m.vnum(f2, m.fv[f][j]) if f2 >= 0 else -1
                f2 = m.ff[f][_prv(j)]
                c.ff[4 * f + j][_prv(j)] = 4 * f2 + m.vnum(f2, m.fv[f][j]) if f2 >= 0 else -1
            for j in range(3):
                a, b = m.fv[f][j], m.fv[f][_nxt(j)]
                ev = edge_vert[(a, b) if a < b else (b, a)]
                c.fv[4 * f + j][j] = a            # the even child keeps the parent's number
                c.fv[4 * f + j][_nxt(j)] = ev
                c.fv[4 * f + _nxt(j)][j] = ev
                c.fv[4 * f + 3][j] = ev
        m = c

    n = len(m.p)
    limit = np.zeros((n, 3), dtype=F32)
    for v in range(n):
        limit[v] = m.weight_boundary(v, F32(1) / F32(5)) if m.boundary[v] else m.weight_one_ring(v, _gamma(m.valence(v)))
    m.p[:] = limit

    normals = np.zeros((n, 3), dtype=F32)
    pi = F32(math.pi)
    for v in range(n):
        ring = [m.p[w] for w in m.one_ring(v)]
        valence = len(ring)
        S = np.zeros(3, dtype=F32)
        T = np.zeros(3, dtype=F32)
        if not m.boundary[v]:
            for j in range(valence):
                ang = F32(F32(F32(F32(2) * pi) * F32(j)) / F32(valence))
                # cosf / sinf of the host's libm: the exact value rounded once (numpy's fp32 routines may be an ulp off)
                S = S + F32(math.cos(float(ang))) * ring[j]
                T = T + F32(math.sin(float(ang))) * ring[j]
        else:
            S = ring[-1] - ring[0]
            if valence == 2:
                T = ring[0] + ring[1] - F32(2) * m.p[v]
            elif valence == 3:
                T = ring[1] - m.p[v]
            elif valence == 4:
                T = F32(-1) * ring[0] + F32(2) * ring[1] + F32(2) * ring[2] + F32(-1) * ring[3] + F32(-2) * m.p[v]
            else:
                theta = F32(pi / F32(valence - 1))
                T = F32(math.sin(float(theta))) * (ring[0] + ring[-1])
                for k in range(1, valence - 1):
                    wt = F32(F32(F32(2) * F32(math.cos(float(theta))) - F32(2)) * F32(math.sin(float(F32(F32(k) * theta)))))
                    T = T + wt * ring[k]
                T = -T
        S = S.astype(F32)
        T = T.astype(F32)
        # Vector3::cross is built on difference_of_products (euclidean_space/vector3.h:151-155): error-compensated, i.e. the
        # exact value rounded once -- fp64 holds the products of fp32 numbers exactly
        S64, T64 = S.astype(np.float64), T.astype(np.float64)
        normals[v] = np.array([S64[1] * T64[2] - S64[2] * T64[1], S64[2] * T64[0] - S64[0] * T64[2], S64[0] * T64[1] - S64[1] * T64[0]]).astype(F32)
    return limit, np.asarray(m.fv, dtype=np.int32).reshape(-1, 3), normals
