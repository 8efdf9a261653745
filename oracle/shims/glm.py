"""Stand-in for PyGLM (`import glm`), TEST INFRASTRUCTURE ONLY.

The reference (`/root/reference/provided/*.py`) does all of its vector maths through
PyGLM, which is not installed in this image and is not pinned by the reference (no
requirements file).  This module restates the documented semantics of the subset the
reference uses: float32 storage for vec3/vec4/mat4, GLSL
definitions of reflect/refract, column-major matrices whose builders post-multiply,
Python-number scalars handled in double.  Summation orders follow GLM's generic
(non-SIMD) C++ so that float32 rounding matches as closely as a restatement can.

PARITY UNPINNED at this boundary: the reference has no tests that pin PyGLM results;
the only external check is the author's stored renders (`renders/*.png`), see
`oracle/make_golden.py`.

Only `oracle/` scripts run in the dev container import this; nothing in the product
package does.
"""
import math
import struct

_F = struct.Struct("f")


def f32(x):
    """Round a Python number to the nearest float32, returned as a Python float."""
    try:
        return _F.unpack(_F.pack(x))[0]
    except OverflowError:
        return math.inf if x > 0 else -math.inf


def _num(x):
    return f32(float(x))


class vec3:
    __slots__ = ("x", "y", "z")
    __array_ufunc__ = None  # numpy scalars defer to __rmul__ etc.

    def __init__(self, *a):
        n = len(a)
        if n == 0:
            self.x = self.y = self.z = 0.0
        elif n == 1:
            v = a[0]
            if isinstance(v, (vec3, vec4)):
                self.x, self.y, self.z = v.x, v.y, v.z
            elif hasattr(v, "__len__"):
                self.x, self.y, self.z = _num(v[0]), _num(v[1]), _num(v[2])
            else:
                self.x = self.y = self.z = _num(v)
        elif n == 3:
            self.x, self.y, self.z = _num(a[0]), _num(a[1]), _num(a[2])
        else:
            raise TypeError("vec3() takes 0, 1 or 3 arguments")

    @staticmethod
    def _raw(x, y, z):
        v = vec3.__new__(vec3)
        v.x, v.y, v.z = x, y, z
        return v

    # container protocol (np.array(vec3), iteration)
    def __len__(self):
        return 3

    def __getitem__(self, i):
        return (self.x, self.y, self.z)[i]

    def __iter__(self):
        return iter((self.x, self.y, self.z))

    def __array__(self, dtype=None, copy=None):
        import numpy as np
        return np.array([self.x, self.y, self.z], dtype=np.float32 if dtype is None else dtype)

    def __repr__(self):
        return f"vec3( {self.x:g}, {self.y:g}, {self.z:g} )"

    def __eq__(self, o):
        if isinstance(o, vec3):
            return self.x == o.x and self.y == o.y and self.z == o.z
        return NotImplemented

    def __ne__(self, o):
        r = self.__eq__(o)
        return r if r is NotImplemented else not r

    __hash__ = None

    def __neg__(self):
        return vec3._raw(-self.x, -self.y, -self.z)

    def __add__(self, o):
        if isinstance(o, vec3):
            return vec3._raw(f32(self.x + o.x), f32(self.y + o.y), f32(self.z + o.z))
        o = _as_vec3_or_num(o)
        if isinstance(o, vec3):
            return self + o
        return vec3._raw(f32(self.x + o), f32(self.y + o), f32(self.z + o))

    __radd__ = __add__

    def __iadd__(self, o):
        r = self + o
        self.x, self.y, self.z = r.x, r.y, r.z
        return self

    def __sub__(self, o):
        if isinstance(o, vec3):
            return vec3._raw(f32(self.x - o.x), f32(self.y - o.y), f32(self.z - o.z))
        o = _as_vec3_or_num(o)
        if isinstance(o, vec3):
            return self - o
        return vec3._raw(f32(self.x - o), f32(self.y - o), f32(self.z - o))

    def __rsub__(self, o):
        o = _as_vec3_or_num(o)
        if isinstance(o, vec3):
            return o - self
        return vec3._raw(f32(o - self.x), f32(o - self.y), f32(o - self.z))

    def __mul__(self, o):
        if isinstance(o, vec3):
            return vec3._raw(f32(self.x * o.x), f32(self.y * o.y), f32(self.z * o.z))
        o = _as_vec3_or_num(o)
        if isinstance(o, vec3):
            return self * o
        return vec3._raw(f32(self.x * o), f32(self.y * o), f32(self.z * o))

    __rmul__ = __mul__

    def __truediv__(self, o):
        if isinstance(o, vec3):
            return vec3._raw(f32(self.x / o.x), f32(self.y / o.y), f32(self.z / o.z))
        o = _num(o)
        return vec3._raw(_div(self.x, o), _div(self.y, o), _div(self.z, o))


def _div(a, b):
    try:
        return f32(a / b)
    except ZeroDivisionError:
        if a == 0 or a != a:
            return math.nan
        return math.copysign(math.inf, a) * math.copysign(1.0, b)


def _as_vec3_or_num(o):
    """Scalars -> float32-rounded Python float; 3-sequences (e.g. ndarray) -> vec3."""
    if hasattr(o, "__len__") and len(o) == 3:
        return vec3(o)
    return _num(o)


class vec4:
    __slots__ = ("x", "y", "z", "w")
    __array_ufunc__ = None

    def __init__(self, *a):
        n = len(a)
        if n == 0:
            self.x = self.y = self.z = self.w = 0.0
        elif n == 1 and isinstance(a[0], vec4):
            v = a[0]
            self.x, self.y, self.z, self.w = v.x, v.y, v.z, v.w
        elif n == 1:
            self.x = self.y = self.z = self.w = _num(a[0])
        elif n == 2 and isinstance(a[0], vec3):
            v = a[0]
            self.x, self.y, self.z, self.w = v.x, v.y, v.z, _num(a[1])
        elif n == 4:
            self.x, self.y, self.z, self.w = (_num(c) for c in a)
        else:
            raise TypeError("unsupported vec4() arguments")

    @staticmethod
    def _raw(x, y, z, w):
        v = vec4.__new__(vec4)
        v.x, v.y, v.z, v.w = x, y, z, w
        return v

    @property
    def xyz(self):
        return vec3._raw(self.x, self.y, self.z)

    def __len__(self):
        return 4

    def __getitem__(self, i):
        return (self.x, self.y, self.z, self.w)[i]

    def __iter__(self):
        return iter((self.x, self.y, self.z, self.w))

    def __repr__(self):
        return f"vec4( {self.x:g}, {self.y:g}, {self.z:g}, {self.w:g} )"

    def __mul__(self, o):
        o = _num(o)
        return vec4._raw(f32(self.x * o), f32(self.y * o), f32(self.z * o), f32(self.w * o))

    __rmul__ = __mul__


class mat4:
    """Column-major 4x4 float32 matrix: self.c[col][row]."""
    __slots__ = ("c",)

    def __init__(self, *a):
        if len(a) == 0:
            d = 1.0
        elif len(a) == 1 and isinstance(a[0], mat4):
            self.c = [list(col) for col in a[0].c]
            return
        elif len(a) == 1:
            d = _num(a[0])
        else:
            raise TypeError("unsupported mat4() arguments")
        self.c = [[d if i == j else 0.0 for i in range(4)] for j in range(4)]

    @staticmethod
    def _cols(cols):
        m = mat4.__new__(mat4)
        m.c = [list(col) for col in cols]
        return m

    def __getitem__(self, j):
        return vec4._raw(*self.c[j])

    def __mul__(self, o):
        if isinstance(o, vec4):
            # GLM generic path: (m0*v0 + m1*v1) + (m2*v2 + m3*v3), per row
            c = self.c
            out = []
            for i in range(4):
                a0 = f32(f32(c[0][i] * o.x) + f32(c[1][i] * o.y))
                a1 = f32(f32(c[2][i] * o.z) + f32(c[3][i] * o.w))
                out.append(f32(a0 + a1))
            return vec4._raw(*out)
        if isinstance(o, mat4):
            # GLM: column j = A0*b[j][0] + A1*b[j][1] + A2*b[j][2] + A3*b[j][3], left to right
            cols = []
            for j in range(4):
                col = []
                for i in range(4):
                    s = f32(self.c[0][i] * o.c[j][0])
                    s = f32(s + f32(self.c[1][i] * o.c[j][1]))
                    s = f32(s + f32(self.c[2][i] * o.c[j][2]))
                    s = f32(s + f32(self.c[3][i] * o.c[j][3]))
                    col.append(s)
                cols.append(col)
            return mat4._cols(cols)
        return NotImplemented

    def __repr__(self):
        return "mat4(" + ", ".join(str(col) for col in self.c) + ")"


# ---------------------------------------------------------------- functions

def dot(a, b):
    if isinstance(a, vec3) and isinstance(b, vec3):
        return f32(f32(f32(a.x * b.x) + f32(a.y * b.y)) + f32(a.z * b.z))
    if isinstance(a, vec4) and isinstance(b, vec4):
        return f32(f32(f32(a.x * b.x) + f32(a.y * b.y)) + f32(f32(a.z * b.z) + f32(a.w * b.w)))
    a, b = _coerce(a), _coerce(b)
    return dot(a, b)


def _coerce(v):
    if isinstance(v, (vec3, vec4)):
        return v
    if hasattr(v, "__len__") and len(v) == 3:
        return vec3(v)
    raise TypeError(f"cannot interpret {type(v)} as a vector")


def length(v):
    v = _coerce(v)
    return f32(math.sqrt(dot(v, v)))


def _inversesqrt(x):
    # GLM: static_cast<T>(1) / sqrt(x) in float
    s = f32(math.sqrt(x)) if x >= 0 else math.nan
    return _div(1.0, s)


def normalize(v):
    v = _coerce(v)
    inv = _inversesqrt(dot(v, v))
    if isinstance(v, vec3):
        return vec3._raw(f32(v.x * inv), f32(v.y * inv), f32(v.z * inv))
    return vec4._raw(f32(v.x * inv), f32(v.y * inv), f32(v.z * inv), f32(v.w * inv))


def cross(a, b):
    return vec3._raw(
        f32(f32(a.y * b.z) - f32(b.y * a.z)),
        f32(f32(a.z * b.x) - f32(b.z * a.x)),
        f32(f32(a.x * b.y) - f32(b.x * a.y)),
    )


def reflect(i, n):
    # GLM: I - N * dot(N, I) * 2
    d = dot(n, i)
    t = n * d
    t = t * 2.0
    return i - t


def refract(i, n, eta):
    eta = _num(eta)
    d = dot(n, i)
    one_minus = f32(1.0 - f32(d * d))
    k = f32(1.0 - f32(f32(eta * eta) * one_minus))
    if k >= 0.0:
        s = f32(f32(eta * d) + f32(math.sqrt(k)))
        return (eta * i) - (s * n)
    return vec3._raw(0.0, 0.0, 0.0)


def radians(x):
    return float(x) * 0.01745329251994329576923690768489


def tan(x):
    return math.tan(float(x))


def translate(m, v):
    c = m.c
    col3 = []
    for i in range(4):
        s = f32(c[0][i] * v.x)
        s = f32(s + f32(c[1][i] * v.y))
        s = f32(s + f32(c[2][i] * v.z))
        s = f32(s + c[3][i])
        col3.append(s)
    return mat4._cols([c[0], c[1], c[2], col3])


def rotate(m, angle, axis):
    a = _num(angle)
    c = f32(math.cos(a))
    s = f32(math.sin(a))
    ax = normalize(axis)
    omc = f32(1.0 - c)
    temp = vec3._raw(f32(omc * ax.x), f32(omc * ax.y), f32(omc * ax.z))
    R = [[0.0] * 3 for _ in range(3)]
    R[0][0] = f32(c + f32(temp.x * ax.x))
    R[0][1] = f32(f32(temp.x * ax.y) + f32(s * ax.z))
    R[0][2] = f32(f32(temp.x * ax.z) - f32(s * ax.y))
    R[1][0] = f32(f32(temp.y * ax.x) - f32(s * ax.z))
    R[1][1] = f32(c + f32(temp.y * ax.y))
    R[1][2] = f32(f32(temp.y * ax.z) + f32(s * ax.x))
    R[2][0] = f32(f32(temp.z * ax.x) + f32(s * ax.y))
    R[2][1] = f32(f32(temp.z * ax.y) - f32(s * ax.x))
    R[2][2] = f32(c + f32(temp.z * ax.z))
    mc = m.c
    cols = []
    for j in range(3):
        col = []
        for i in range(4):
            v = f32(mc[0][i] * R[j][0])
            v = f32(v + f32(mc[1][i] * R[j][1]))
            v = f32(v + f32(mc[2][i] * R[j][2]))
            col.append(v)
        cols.append(col)
    cols.append(list(mc[3]))
    return mat4._cols(cols)


def scale(m, v):
    c = m.c
    return mat4._cols([
        [f32(c[0][i] * v.x) for i in range(4)],
        [f32(c[1][i] * v.y) for i in range(4)],
        [f32(c[2][i] * v.z) for i in range(4)],
        list(c[3]),
    ])


def transpose(m):
    return mat4._cols([[m.c[i][j] for i in range(4)] for j in range(4)])


def inverse(m):
    """GLM compute_inverse<4,4>: cofactor expansion, float32 at every step."""
    c = m.c

    def M(col, row):
        return c[col][row]

    def sub2(a, b, cc, d):  # a*b - cc*d in float32
        return f32(f32(a * b) - f32(cc * d))

    Coef00 = sub2(M(2, 2), M(3, 3), M(3, 2), M(2, 3))
    Coef02 = sub2(M(1, 2), M(3, 3), M(3, 2), M(1, 3))
    Coef03 = sub2(M(1, 2), M(2, 3), M(2, 2), M(1, 3))
    Coef04 = sub2(M(2, 1), M(3, 3), M(3, 1), M(2, 3))
    Coef06 = sub2(M(1, 1), M(3, 3), M(3, 1), M(1, 3))
    Coef07 = sub2(M(1, 1), M(2, 3), M(2, 1), M(1, 3))
    Coef08 = sub2(M(2, 1), M(3, 2), M(3, 1), M(2, 2))
    Coef10 = sub2(M(1, 1), M(3, 2), M(3, 1), M(1, 2))
    Coef11 = sub2(M(1, 1), M(2, 2), M(2, 1), M(1, 2))
    Coef12 = sub2(M(2, 0), M(3, 3), M(3, 0), M(2, 3))
    Coef14 = sub2(M(1, 0), M(3, 3), M(3, 0), M(1, 3))
    Coef15 = sub2(M(1, 0), M(2, 3), M(2, 0), M(1, 3))
    Coef16 = sub2(M(2, 0), M(3, 2), M(3, 0), M(2, 2))
    Coef18 = sub2(M(1, 0), M(3, 2), M(3, 0), M(1, 2))
    Coef19 = sub2(M(1, 0), M(2, 2), M(2, 0), M(1, 2))
    Coef20 = sub2(M(2, 0), M(3, 1), M(3, 0), M(2, 1))
    Coef22 = sub2(M(1, 0), M(3, 1), M(3, 0), M(1, 1))
    Coef23 = sub2(M(1, 0), M(2, 1), M(2, 0), M(1, 1))

    Fac0 = (Coef00, Coef00, Coef02, Coef03)
    Fac1 = (Coef04, Coef04, Coef06, Coef07)
    Fac2 = (Coef08, Coef08, Coef10, Coef11)
    Fac3 = (Coef12, Coef12, Coef14, Coef15)
    Fac4 = (Coef16, Coef16, Coef18, Coef19)
    Fac5 = (Coef20, Coef20, Coef22, Coef23)

    Vec0 = (M(1, 0), M(0, 0), M(0, 0), M(0, 0))
    Vec1 = (M(1, 1), M(0, 1), M(0, 1), M(0, 1))
    Vec2 = (M(1, 2), M(0, 2), M(0, 2), M(0, 2))
    Vec3 = (M(1, 3), M(0, 3), M(0, 3), M(0, 3))

    def comb(Va, Fa, Vb, Fb, Vc, Fc):  # Va*Fa - Vb*Fb + Vc*Fc
        return [f32(f32(f32(Va[k] * Fa[k]) - f32(Vb[k] * Fb[k])) + f32(Vc[k] * Fc[k])) for k in range(4)]

    Inv0 = comb(Vec1, Fac0, Vec2, Fac1, Vec3, Fac2)
    Inv1 = comb(Vec0, Fac0, Vec2, Fac3, Vec3, Fac4)
    Inv2 = comb(Vec0, Fac1, Vec1, Fac3, Vec3, Fac5)
    Inv3 = comb(Vec0, Fac2, Vec1, Fac4, Vec2, Fac5)

    SignA = (+1.0, -1.0, +1.0, -1.0)
    SignB = (-1.0, +1.0, -1.0, +1.0)
    Inverse = [
        [Inv0[k] * SignA[k] for k in range(4)],
        [Inv1[k] * SignB[k] for k in range(4)],
        [Inv2[k] * SignA[k] for k in range(4)],
        [Inv3[k] * SignB[k] for k in range(4)],
    ]
    Row0 = (Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0])
    Dot0 = [f32(c[0][k] * Row0[k]) for k in range(4)]
    Dot1 = f32(f32(Dot0[0] + Dot0[1]) + f32(Dot0[2] + Dot0[3]))
    OneOverDet = _div(1.0, Dot1)
    return mat4._cols([[f32(Inverse[j][i] * OneOverDet) for i in range(4)] for j in range(4)])
