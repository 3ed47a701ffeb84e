"""A 60-line stand-in for the few `gtsam` classes the reference's CustomFactor callbacks touch
(perseus/smoother/factors.py:8-275): TEST INFRASTRUCTURE ONLY.  gtsam itself is not installable in this image
(SURVEY.md §8(c)); this stub lets tests drive `error_func(this, v, H)` exactly the way GTSAM's optimiser does --
one factor at a time, the same Values for every factor of a linearisation, Jacobians handed over as a list that the
callback assigns into -- so that perseus_b200/smoother/_gtsam_factors.py is exercised.  It implements no geometry."""
import numpy as np


class Pose3:
    def __init__(self, *args):
        if len(args) == 0:
            self._m = np.eye(4)
        elif len(args) == 1:
            self._m = np.array(args[0], dtype=np.float64).reshape(4, 4)
        else:  # (Rot3 | 3x3, translation)
            R = args[0].matrix() if hasattr(args[0], "matrix") else np.asarray(args[0], dtype=np.float64)
            self._m = np.eye(4)
            self._m[:3, :3] = R
            self._m[:3, 3] = np.asarray(args[1], dtype=np.float64)

    def matrix(self):
        return self._m.copy()

    def rotation(self):
        return Rot3(self._m[:3, :3])

    def translation(self):
        return self._m[:3, 3].copy()


class Rot3:
    def __init__(self, R=None):
        self._R = np.eye(3) if R is None else np.array(R, dtype=np.float64).reshape(3, 3)

    def matrix(self):
        return self._R.copy()


class Cal3_S2:
    def __init__(self, fx=1.0, fy=1.0, s=0.0, u0=0.0, v0=0.0):
        self._k = (float(fx), float(fy), float(s), float(u0), float(v0))

    def fx(self): return self._k[0]  # noqa: E704
    def fy(self): return self._k[1]  # noqa: E704
    def skew(self): return self._k[2]  # noqa: E704
    def px(self): return self._k[3]  # noqa: E704
    def py(self): return self._k[4]  # noqa: E704


class Values:
    def __init__(self):
        self._d = {}

    def insert(self, key, value):
        self._d[int(key)] = value

    def update(self, key, value):
        self._d[int(key)] = value

    def exists(self, key):
        return int(key) in self._d

    def _at(self, key):
        if int(key) not in self._d:
            raise RuntimeError(f"ValuesKeyDoesNotExist: key {key}")  # gtsam raises a RuntimeError subclass
        return self._d[int(key)]

    def atPose3(self, key):
        return self._at(key)

    def atVector(self, key):
        return np.array(self._at(key), dtype=np.float64)


class CustomFactor:
    """gtsam.CustomFactor(noiseModel, keys, error_func): the optimiser calls error_func(self, values, H) with
    H = None (error only) or a list of len(keys) matrices to be overwritten."""

    def __init__(self, noise_model, keys, error_func):
        self._noise, self._keys, self._error_func = noise_model, [int(k) for k in keys], error_func

    def keys(self):
        return list(self._keys)

    def unwhitenedError(self, values, H=None):
        return self._error_func(self, values, H)


class _Diagonal:
    @staticmethod
    def Sigmas(s):
        return ("diagonal", np.asarray(s, dtype=np.float64))


class noiseModel:  # noqa: N801  (gtsam's own spelling)
    Diagonal = _Diagonal
