"""Stand-in for the libigl Python bindings (`import igl`), TEST INFRASTRUCTURE ONLY.

The reference uses two libigl entry points (`/root/reference/provided/geometry/mesh.py:20,108`);
libigl is not installed here and not pinned by the reference.  Restated:

* `read_obj(path)` -> (V, TC, N, F, FTC, FN): `v`, `vt`, `vn` rows as float64 arrays and
  0-based integer faces (polygons are fan-triangulated, as libigl's reader does for the
  triangle-mesh overload).
* `barycentric_coordinates_tri(P, A, B, C)`: libigl's `barycentric_coordinates` for triangles
  (the dot-product formulation: v0=B-A, v1=C-A, v2=P-A; d00,d01,d11,d20,d21;
  l1=(d11*d20-d01*d21)/den, l2=(d00*d21-d01*d20)/den, l0=1-(l1+l2)), evaluated in the
  dtype of the inputs (the reference passes float32 rows) and squeezed to shape (3,)
  for a single query row, which is what `mesh.py:109-111` indexes.

PARITY UNPINNED at this boundary (no reference test pins libigl output); cross-checked
softly by `renders/TorusMesh.png` in `oracle/make_golden.py`.
"""
import numpy as np


def read_obj(path):
    V, TC, N, F, FTC, FN = [], [], [], [], [], []
    with open(path) as fh:
        for line in fh:
            parts = line.split()
            if not parts:
                continue
            tag = parts[0]
            if tag == "v":
                V.append([float(x) for x in parts[1:4]])
            elif tag == "vt":
                TC.append([float(x) for x in parts[1:3]])
            elif tag == "vn":
                N.append([float(x) for x in parts[1:4]])
            elif tag == "f":
                vi, ti, ni = [], [], []
                for tok in parts[1:]:
                    sp = tok.split("/")
                    vi.append(int(sp[0]) - 1 if int(sp[0]) > 0 else len(V) + int(sp[0]))
                    if len(sp) > 1 and sp[1]:
                        ti.append(int(sp[1]) - 1)
                    if len(sp) > 2 and sp[2]:
                        ni.append(int(sp[2]) - 1)
                for k in range(1, len(vi) - 1):
                    F.append([vi[0], vi[k], vi[k + 1]])
                    if len(ti) == len(vi):
                        FTC.append([ti[0], ti[k], ti[k + 1]])
                    if len(ni) == len(vi):
                        FN.append([ni[0], ni[k], ni[k + 1]])
    f64 = lambda a, w: np.array(a, dtype=np.float64).reshape(-1, w)
    i64 = lambda a: np.array(a, dtype=np.int64).reshape(-1, 3)
    return f64(V, 3), f64(TC, 2), f64(N, 3), i64(F), i64(FTC), i64(FN)


def barycentric_coordinates_tri(P, A, B, C):
    P, A, B, C = (np.asarray(x) for x in (P, A, B, C))
    dt = P.dtype if P.dtype in (np.float32, np.float64) else np.float64
    P, A, B, C = (x.astype(dt).reshape(-1, 3) for x in (P, A, B, C))
    v0 = B - A
    v1 = C - A
    v2 = P - A

    def rowsum(x):  # Eigen rowwise().sum() on 3 columns: (x0 + x1) + x2
        return (x[:, 0] + x[:, 1]) + x[:, 2]

    d00 = rowsum(v0 * v0)
    d01 = rowsum(v0 * v1)
    d11 = rowsum(v1 * v1)
    d20 = rowsum(v2 * v0)
    d21 = rowsum(v2 * v1)
    with np.errstate(all="ignore"):
        denom = d00 * d11 - d01 * d01
        L = np.empty((P.shape[0], 3), dtype=dt)
        L[:, 1] = (d11 * d20 - d01 * d21) / denom
        L[:, 2] = (d00 * d21 - d01 * d20) / denom
        L[:, 0] = dt.type(1.0) - (L[:, 1] + L[:, 2])
    return L[0] if L.shape[0] == 1 else L
