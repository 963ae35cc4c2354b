"""Deterministic ray batches for seam-B2 parity: P primary rays, U incoherent rays, D diffuse
bounce rays.  Pure numpy, seeded, so the oracle and the CUDA core read identical bits."""
import numpy as np

RAY_DT = np.dtype([("origin", "<f4", 3), ("t_min", "<f4"), ("direction", "<f4", 3), ("t_max", "<f4")])
FMAX = np.float32(3.4028234663852886e38)


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def scene_bounds(fs):
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    for inst in fs.instances:
        x = np.asarray(inst["xform"], np.float64)
        m, t = x[:9].reshape(3, 3).T, x[9:]
        if inst["prim_kind"] == 0:
            s = fs.spheres[inst["prim_index"]]
            c, r = np.asarray(s["center"]), s["radius"]
            plo, phi = c - r, c + r
        else:
            p = fs.meshes[inst["prim_index"]]["positions"]
            plo, phi = p.min(0), p.max(0)
        for k in range(8):
            q = np.array([phi[0] if k & 4 else plo[0], phi[1] if k & 2 else plo[1], phi[2] if k & 1 else plo[2]])
            w = m @ q + t
            lo, hi = np.minimum(lo, w), np.maximum(hi, w)
    return lo, hi


def primary_rays(fs, width, height, camera=0):
    c = fs.cameras[camera]
    eye, f, up = (np.asarray(c[k], np.float64) for k in ("eye", "forward", "up"))
    f = f / np.linalg.norm(f)
    r = np.cross(f, up); r /= np.linalg.norm(r)
    u = np.cross(r, f)
    hc = 0.5 / np.tan(c["fov"] * 0.5)
    aspect = width / height
    j, i = np.mgrid[0:height, 0:width]
    x = ((i + 0.5) / width - 0.5) * aspect
    y = ((height - j - 1) + 0.5) / height - 0.5
    d = f * hc + r * x[..., None] + u * y[..., None]
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    rays = np.zeros(width * height, RAY_DT)
    rays["origin"] = eye.astype(np.float32)
    rays["direction"] = d.reshape(-1, 3).astype(np.float32)
    rays["t_min"] = 1e-4
    rays["t_max"] = FMAX
    return rays


def incoherent_rays(fs, n, seed=0x5EED0003, finite_tmax=False):
    g = _rng(seed)
    lo, hi = scene_bounds(fs)
    c, h = 0.5 * (lo + hi), 0.75 * (hi - lo)
    rays = np.zeros(n, RAY_DT)
    rays["origin"] = (c + (g.random((n, 3)) * 2 - 1) * h).astype(np.float32)
    z = g.random(n) * 2 - 1
    ph = g.random(n) * 2 * np.pi
    s = np.sqrt(1 - z * z)
    rays["direction"] = np.stack([s * np.cos(ph), s * np.sin(ph), z], -1).astype(np.float32)
    rays["t_min"] = 1e-4
    rays["t_max"] = (g.random(n) * np.linalg.norm(hi - lo)).astype(np.float32) if finite_tmax else FMAX
    return rays


def bounce_rays(rays, hits, seed=0x5EED0002):
    """Cosine-hemisphere bounce rays from the hit points of `rays` (hemisphere oriented against the incoming ray).
    t_min follows pt.rs:184."""
    g = _rng(seed)
    ok = hits["inst_id"] >= 0
    o = rays["origin"][ok].astype(np.float64) + rays["direction"][ok].astype(np.float64) * hits["t"][ok][:, None]
    n = -rays["direction"][ok].astype(np.float64)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    m = len(o)
    u0, u1 = g.random(m), g.random(m)
    ph = 2 * np.pi * u0
    st, ct = np.sqrt(u1), np.sqrt(1 - u1)
    a = np.where(np.abs(n[:, :1]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    t = np.cross(n, a); t /= np.linalg.norm(t, axis=-1, keepdims=True)
    b = np.cross(n, t)
    d = t * (st * np.cos(ph))[:, None] + b * (st * np.sin(ph))[:, None] + n * ct[:, None]
    out = np.zeros(m, RAY_DT)
    out["origin"] = o.astype(np.float32)
    out["direction"] = d.astype(np.float32)
    out["t_min"] = (1e-4 / np.maximum(ct, 1e-5)).astype(np.float32)
    out["t_max"] = FMAX
    return out


def compare_hits(ref, got, rel=1e-5):
    """ids equal, or an equal-t tie; |dt|/max(t,1e-6) <= rel on matches.  Returns a dict of rates."""
    n = len(ref)
    both_miss = (ref["inst_id"] < 0) & (got["inst_id"] < 0)
    both_hit = (ref["inst_id"] >= 0) & (got["inst_id"] >= 0)
    rt, gt = ref["t"].astype(np.float64), got["t"].astype(np.float64)
    dt = np.abs(rt - gt) / np.maximum(np.abs(rt), 1e-6)
    same_id = both_hit & (ref["inst_id"] == got["inst_id"]) & (ref["prim_id"] == got["prim_id"])
    tie = both_hit & ~same_id & (dt <= rel)
    match = both_miss | same_id | tie
    t_ok = ~same_id | (dt <= rel)
    return dict(n=n, match_rate=float(match.mean()) if n else 1.0, tie_rate=float(tie.mean()) if n else 0.0,
                t_ok_rate=float(t_ok.mean()) if n else 1.0, max_dt_same_id=float(dt[same_id].max()) if same_id.any() else 0.0,
                hit_rate=float((ref["inst_id"] >= 0).mean()) if n else 0.0,
                bit_exact_t=float((ref["t"][same_id] == got["t"][same_id]).mean()) if same_id.any() else 1.0)
