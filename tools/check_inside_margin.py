"""Monte-Carlo check of the margin in pt_trace.cuh's PT_INSIDE_APPROX variant (the area-sum inside test with MUFU.RSQ in
front of the exact one): fp32 emulation of the reference's staged test (triangles.h:56-70) against the approximate
decision with the reciprocal square root perturbed by up to 2^-22 either way, on hit points placed AT the acceptance
threshold of triangles from 1e-3 to 20 units.  A confident decision must never contradict the exact test.
60 M cases: 0 contradictions, 23 % ambiguous (those take the exact path; ordinary rays are almost never that close).
Second part, PT_TRI_PREREJECT (the early exit in tri_intersect_ref from an approximate quotient): 20 M threshold cases,
none of the 8.4 M candidates the exact test accepts is rejected early, 7.7 M leave early."""
import numpy as np

f32 = np.float32


def cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - b[:, 1] * a[:, 2], a[:, 2] * b[:, 0] - b[:, 2] * a[:, 0],
                     a[:, 0] * b[:, 1] - b[:, 0] * a[:, 1]], 1).astype(f32)


def dot(a, b):
    return ((a[:, 0] * b[:, 0]).astype(f32) + (a[:, 1] * b[:, 1]).astype(f32)).astype(f32) + (a[:, 2] * b[:, 2]).astype(f32)


def main(reps=20, n=200000):
    rng = np.random.default_rng(0)
    eps = f32(1e-4)
    bad = amb = tot = 0
    for scale in (1e-3, 1e-2, 0.1, 1, 20):
        for _ in range(reps):
            A = rng.uniform(-10, 10, (n, 3)).astype(f32)
            B = (A + rng.normal(size=(n, 3)) * scale).astype(f32)
            C = (A + rng.normal(size=(n, 3)) * scale).astype(f32)
            nrm = cross((B - A).astype(f32), (C - A).astype(f32))
            S = np.sqrt(dot(nrm, nrm)).astype(f32)
            u = rng.random((n, 1)).astype(f32)
            E = (A + (B - A) * u).astype(f32)
            L = np.linalg.norm((B - A).astype(np.float64), axis=1)
            out = np.cross(nrm.astype(np.float64), (B - A).astype(np.float64))
            out /= np.linalg.norm(out, axis=1, keepdims=True) + 1e-300
            k = rng.choice([0.5, 0.9, 0.99, 0.999, 1.0, 1.001, 1.01, 1.1, 2.0, -1.0], (n, 1))
            h = (eps / np.maximum(L, 1e-30))[:, None] * k  # area excess L * h = k * eps
            P = (E.astype(np.float64) + out * h * rng.choice([-1, 1], (n, 1))).astype(f32)
            f0, f1, f2 = (P - A).astype(f32), (P - B).astype(f32), (P - C).astype(f32)
            c1, c2, c3 = cross(f0, f1), cross(f0, f2), cross(f2, f1)
            q1, q2, q3 = dot(c1, c1), dot(c2, c2), dot(c3, c3)
            s1, s2, s3 = np.sqrt(q1).astype(f32), np.sqrt(q2).astype(f32), np.sqrt(q3).astype(f32)
            Se = (S + eps).astype(f32)
            exact = ~(s1 > Se) & ~((s1 + s2).astype(f32) > Se) & \
                ~(np.abs(((S - s1).astype(f32) - s2).astype(f32) - s3).astype(f32) > eps)
            with np.errstate(all="ignore"):
                for sgn in (-1, 1, 0):
                    def rs(q):
                        return ((1 / np.sqrt(q.astype(np.float64))) * (1 + sgn * 2.0 ** -22 * rng.random(n))).astype(f32)
                    sm = (((q1 * rs(q1)).astype(f32) + (q2 * rs(q2)).astype(f32)).astype(f32) + (q3 * rs(q3)).astype(f32)).astype(f32)
                    dev = np.abs((S - sm).astype(f32))
                    margin = (f32(8e-6) * (S + sm).astype(f32)).astype(f32)
                    rej, acc = dev > (eps + margin).astype(f32), dev < (eps - margin).astype(f32)
                    bad += int((rej & exact).sum() + (acc & ~exact).sum())
                    amb += int((~rej & ~acc).sum())
                    tot += n
    print(f"confident contradictions: {bad}; ambiguous (exact path): {amb} of {tot} ({amb / tot:.2%})")
    return bad


def prereject(reps=10, n=200000):
    """PT_TRI_PREREJECT: the same sum from a hit point that comes from an approximate quotient; a reject must never
    hit a candidate the exact test (IEEE quotient, IEEE square roots) accepts."""
    rng = np.random.default_rng(1)
    eps = f32(1e-4)
    bad = rej_n = acc_n = tot = 0
    for scale in (1e-3, 1e-2, 0.1, 1, 20):
        for _ in range(reps):
            A = rng.uniform(-10, 10, (n, 3)).astype(f32)
            B = (A + rng.normal(size=(n, 3)) * scale).astype(f32)
            C = (A + rng.normal(size=(n, 3)) * scale).astype(f32)
            nrm = cross((B - A).astype(f32), (C - A).astype(f32))
            S = np.sqrt(dot(nrm, nrm)).astype(f32)
            nu = (nrm * (f32(1) / S)[:, None]).astype(f32)
            w = (-dot(nu, A)).astype(f32)
            u = rng.random((n, 1)).astype(f32)
            E = (A + (B - A) * u).astype(f32)
            L = np.linalg.norm((B - A).astype(np.float64), axis=1)
            out = np.cross(nrm.astype(np.float64), (B - A).astype(np.float64))
            out /= np.linalg.norm(out, axis=1, keepdims=True) + 1e-300
            k = rng.choice([0.0, 0.5, 0.9, 0.99, 1.0, 1.01, 1.1, 2.0, 10.0, -1.0], (n, 1))
            h = (eps / np.maximum(L, 1e-30))[:, None] * k
            target = E.astype(np.float64) + out * h * rng.choice([-1, 1], (n, 1))
            o = rng.uniform(-20, 20, (n, 3)).astype(f32)
            dd = target - o
            d = (dd / (np.linalg.norm(dd, axis=1, keepdims=True) + 1e-300)).astype(f32)
            with np.errstate(all="ignore"):
                sd = dot(d, nu)
                num = (-(dot(o, nu) + w).astype(f32)).astype(f32)
                nd = (num / sd).astype(f32)

                def areas(t, perturb):
                    P = (o + (d * t[:, None]).astype(f32)).astype(f32)
                    f0, f1, f2 = (P - A).astype(f32), (P - B).astype(f32), (P - C).astype(f32)
                    c1, c2, c3 = cross(f0, f1), cross(f0, f2), cross(f2, f1)
                    return (f0, f1, f2), (dot(c1, c1), dot(c2, c2), dot(c3, c3))
                _, (q1, q2, q3) = areas(nd, 0)
                s1, s2, s3 = np.sqrt(q1).astype(f32), np.sqrt(q2).astype(f32), np.sqrt(q3).astype(f32)
                Se = (S + eps).astype(f32)
                exact = ~(s1 > Se) & ~((s1 + s2).astype(f32) > Se) & \
                    ~(np.abs(((S - s1).astype(f32) - s2).astype(f32) - s3).astype(f32) > eps)
                exact &= (nd >= eps) & np.isfinite(nd)
                for sgn in (-1, 1):
                    tq = (num * ((1 / sd.astype(np.float64)) * (1 + sgn * 2.0 ** -22.5 * rng.random(n))).astype(f32)).astype(f32)
                    (g0, g1, g2), (p1, p2, p3) = areas(tq, 0)

                    def rs(q):
                        return ((1 / np.sqrt(q.astype(np.float64))) * (1 - sgn * 2.0 ** -22 * rng.random(n))).astype(f32)
                    sm = (((p1 * rs(p1)).astype(f32) + (p2 * rs(p2)).astype(f32)).astype(f32) + (p3 * rs(p3)).astype(f32)).astype(f32)
                    l1 = (np.abs(g0).sum(1) + np.abs(g1).sum(1) + np.abs(g2).sum(1)).astype(f32)
                    dP = (f32(1e-6) * (np.abs(tq) + np.abs(o).sum(1)).astype(f32)).astype(f32)
                    margin = ((f32(8e-6) * (S + sm).astype(f32)).astype(f32) + (f32(2) * l1 * dP).astype(f32) +
                              (f32(1e-6) * l1 * l1).astype(f32)).astype(f32)
                    rej = np.abs((S - sm).astype(f32)) > (eps + margin).astype(f32)
                    bad += int((rej & exact).sum())
                    rej_n += int(rej.sum())
                    acc_n += int(exact.sum())
                    tot += n
    print(f"pre-reject: {bad} accepted candidates rejected; {rej_n} of {tot} rejected early, {acc_n} accepted by the exact test")
    return bad


if __name__ == "__main__":
    raise SystemExit(1 if main() + prereject() else 0)
