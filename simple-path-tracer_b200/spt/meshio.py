"""Mesh ingestion for the host side: a small OBJ reader with the semantics of `tobj::load_obj`
(`triangulate = true`, `single_index = true`) as used by TriMesh::load (src/primitive/triangle.rs:57-108),
plus TriMesh::calc_tangents (triangle.rs:339-388).  `.npz` meshes (positions/normals/texcoords/indices)
are accepted too, so converted assets can travel without the OBJ text.
"""
import numpy as np

f32 = np.float32


def load_obj(path):
    """Returns (positions[n,3], normals[n,3] or None, texcoords[n,2] or None, indices[t,3]) as float32/uint32.

    All `o`/`g` models of the file are merged the way TriMesh::load does (triangle.rs:57-108), except for its index-offset
    quirk Q22 (offset = number of indices loaded so far, triangle.rs:96-102), which is fixed: offset = running vertex count."""
    v, vt, vn = [], [], []
    models = [[]]  # list of face lists; each face = list of (vi, ti, ni)
    with open(path) as fh:
        for line in fh:
            s = line.split()
            if not s:
                continue
            k = s[0]
            if k == "v":
                v.append([float(x) for x in s[1:4]])
            elif k == "vt":
                vt.append([float(x) for x in s[1:3]])
            elif k == "vn":
                vn.append([float(x) for x in s[1:4]])
            elif k in ("o", "g"):
                if models[-1]:
                    models.append([])
            elif k == "f":
                face = []
                for tok in s[1:]:
                    parts = tok.split("/")
                    vi = int(parts[0])
                    ti = int(parts[1]) if len(parts) > 1 and parts[1] else 0
                    ni = int(parts[2]) if len(parts) > 2 and parts[2] else 0
                    vi = vi - 1 if vi > 0 else len(v) + vi
                    ti = (ti - 1 if ti > 0 else len(vt) + ti) if ti != 0 else -1
                    ni = (ni - 1 if ni > 0 else len(vn) + ni) if ni != 0 else -1
                    face.append((vi, ti, ni))
                models[-1].append(face)
    pos_all, nrm_all, uv_all, idx_all = [], [], [], []
    have_n = have_t = False
    n_vertices = 0
    for faces in models:
        if not faces:
            continue
        remap = {}
        mpos, mnrm, muv, midx = [], [], [], []
        m_has_n = all(c[2] >= 0 for f in faces for c in f)
        m_has_t = all(c[1] >= 0 for f in faces for c in f)
        for f in faces:
            corners = []
            for c in f:
                key = c
                if key not in remap:
                    remap[key] = len(mpos)
                    mpos.append(v[c[0]])
                    mnrm.append(vn[c[2]] if m_has_n else [0.0, 0.0, 1.0])  # MeshVertex::default, triangle.rs:29-39
                    muv.append(vt[c[1]] if m_has_t else [0.0, 0.0])
                corners.append(remap[key])
            for k in range(1, len(corners) - 1):  # fan triangulation
                midx.append([corners[0], corners[k], corners[k + 1]])
        have_n |= m_has_n
        have_t |= m_has_t
        pos_all += mpos
        nrm_all += mnrm
        uv_all += muv
        # Q22 FIXED: the reference offsets by the number of INDICES loaded so far (triangle.rs:96-102), which panics or wires
        # triangles to the wrong vertices for any OBJ with more than one object; the evident intent is the running vertex count
        idx_all += [[a + n_vertices, b + n_vertices, c + n_vertices] for a, b, c in midx]
        n_vertices += len(mpos)
    return (np.asarray(pos_all, f32).reshape(-1, 3), np.asarray(nrm_all, f32).reshape(-1, 3),
            np.asarray(uv_all, f32).reshape(-1, 2), np.asarray(idx_all, np.uint32).reshape(-1, 3))


def load_mesh(path):
    if str(path).endswith(".npz"):
        z = np.load(path)
        pos = z["positions"].astype(f32)
        n = len(pos)
        nrm = z["normals"].astype(f32) if "normals" in z else np.tile(np.array([0, 0, 1], f32), (n, 1))
        uv = z["texcoords"].astype(f32) if "texcoords" in z else np.zeros((n, 2), f32)
        return pos, nrm, uv, z["indices"].astype(np.uint32).reshape(-1, 3)
    return load_obj(path)


def _normalize_rows(a):
    ln = np.sqrt((a[:, 0] * a[:, 0] + a[:, 1] * a[:, 1]) + a[:, 2] * a[:, 2], dtype=f32)
    with np.errstate(divide="ignore", invalid="ignore"):
        return (a / ln[:, None]).astype(f32)


def calc_tangents(pos, uv, idx):
    """TriMesh::calc_tangents, triangle.rs:339-388.  Vertices no UV-nondegenerate triangle touches keep the
    MeshVertex defaults tangent = X, bitangent = Y."""
    n = len(pos)
    tan = np.tile(np.array([1, 0, 0], f32), (n, 1))
    bit = np.tile(np.array([0, 1, 0], f32), (n, 1))
    if len(idx) == 0:
        return tan, bit
    i0, i1, i2 = idx[:, 0], idx[:, 1], idx[:, 2]
    e1 = pos[i1] - pos[i0]
    e2 = pos[i2] - pos[i0]
    u1 = uv[i1] - uv[i0]
    u2 = uv[i2] - uv[i0]
    det = (u1[:, 0] * u2[:, 1] - u1[:, 1] * u2[:, 0]).astype(f32)
    ok = det != 0
    if not ok.any():
        return tan, bit
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (f32(1.0) / det).astype(f32)
    t = _normalize_rows(((e1 * u2[:, 1:2] - e2 * u1[:, 1:2]) * inv[:, None]).astype(f32))
    b = _normalize_rows(((e2 * u1[:, 0:1] - e1 * u2[:, 0:1]) * inv[:, None]).astype(f32))
    tsum = np.zeros((n, 3), f32)
    bsum = np.zeros((n, 3), f32)
    deg = np.zeros(n, np.int64)
    sel = np.nonzero(ok)[0]
    # sequential accumulation in triangle order (np.add.at is unbuffered and ordered)
    order = np.stack([i0[sel], i1[sel], i2[sel]], axis=1).reshape(-1)
    np.add.at(tsum, order, np.repeat(t[sel], 3, axis=0))
    np.add.at(bsum, order, np.repeat(b[sel], 3, axis=0))
    np.add.at(deg, order, 1)
    has = deg != 0
    inv_deg = (f32(1.0) / deg[has].astype(f32)).astype(f32)
    tan[has] = _normalize_rows((tsum[has] * inv_deg[:, None]).astype(f32))
    bit[has] = _normalize_rows((bsum[has] * inv_deg[:, None]).astype(f32))
    return tan, bit
