"""Host-side mirror of the reference's operator interface for the hot path.

Same names, argument meaning and error behaviour as the hpp::core classes, each backed by the
C ABI (no arithmetic here).  Single-item calls keep the reference signature
`validate(q) -> (bool, report)` / `validate(path, reverse) -> (bool, validPart, report)`;
`validate_batch` is the batch extension the engine exists for.

  CollisionValidation      reference include/hpp/core/collision-validation.hh, src/collision-validation.cc:48-95
  JointBoundValidation     reference src/joint-bound-validation.cc:45-80
  ConfigValidations        reference src/config-validations.cc:40-49
  Discretized              reference src/path-validation/discretized.cc:55-107
  Progressive / Dichotomy  reference src/continuous-validation/{progressive,dichotomy}.cc
  StraightPath             reference src/straight-path.cc:91-158 (value object only)
"""
import numpy as np


class ValidationReport:
    """CollisionValidationReport / JointBoundValidationReport / PathValidationReport rolled into a
    plain record (reference include/hpp/core/collision-validation-report.hh:47-107,
    include/hpp/core/path-validation-report.hh:45-70)."""

    def __init__(self, reason, index, parameter=None, model=None):
        self.reason = int(reason)
        self.index = int(index)          # colliding pair index / violated configuration rank
        self.parameter = parameter       # PathValidationReport::parameter
        self.object1 = self.object2 = None
        if model is not None and self.reason == 1 and self.index >= 0:
            self.pair_index = self.index

    def __repr__(self):
        kind = {0: "valid", 1: "collision", 2: "joint bound", 3: "iteration cap"}.get(self.reason, "?")
        return "ValidationReport(%s, index=%d, parameter=%r)" % (kind, self.index, self.parameter)


class StraightPath:
    """Edge value object: StraightPath::create(robot, q1, q2, length) over [0, length]."""

    def __init__(self, q1, q2, length):
        self.initial = np.asarray(q1, np.float64)
        self.end = np.asarray(q2, np.float64)
        self.length = float(length)

    def timeRange(self):
        return (0.0, self.length)

    def extract(self, t0, t1, workspace=None):
        """StraightPath::impl_extract (src/straight-path.cc:142-158): a new path between eval(t0) and eval(t1).
        With a workspace the two configurations come from hppb200_path_eval (any joint type); without one the
        interpolation is done here and is only right for vector-space joints."""
        if workspace is not None:
            q = workspace.path_eval([self.initial, self.initial], [self.end, self.end], [self.length] * 2, [t0, t1])
            return StraightPath(q[0], q[1], abs(t1 - t0))
        def at(t):
            if t == 0.0 or self.length == 0.0:
                return self.initial.copy()
            if t == self.length:
                return self.end.copy()
            u = t / self.length
            return self.initial + u * (self.end - self.initial)
        return StraightPath(at(t0), at(t1), abs(t1 - t0))


class _ConfigValidator:
    _check_bounds = False
    _collision = True

    def __init__(self, workspace):
        self.ws = workspace

    def validate(self, q):
        """bool validate(const Configuration_t&, ValidationReportPtr_t&)"""
        ok, rep = self.validate_batch(np.asarray(q, np.float64).reshape(1, -1))
        return bool(ok[0]), rep[0]

    def validate_batch(self, q):
        valid, first, reason = self.ws.validate_configs(q, check_bounds=self._check_bounds)
        reports = [None if v else ValidationReport(r, f) for v, f, r in zip(valid, first, reason)]
        return valid.astype(bool), reports


class CollisionValidation(_ConfigValidator):
    @classmethod
    def create(cls, workspace):
        return cls(workspace)


class ConfigValidations(_ConfigValidator):
    """JointBoundValidation then CollisionValidation, first failure wins (default Problem)."""
    _check_bounds = True

    @classmethod
    def create(cls, workspace):
        return cls(workspace)


class JointBoundValidation:
    """Bounds only.  Host-side compare (nq comparisons per configuration, setup-scale work);
    the batched device path checks the same predicate inside ConfigValidations."""

    def __init__(self, workspace):
        self.ws = workspace
        d = workspace.model.desc
        self.lower, self.upper = d.q_lower, d.q_upper

    @classmethod
    def create(cls, workspace):
        return cls(workspace)

    def validate(self, q):
        q = np.asarray(q, np.float64)
        for i in range(q.shape[0]):
            if self.lower[i] > self.upper[i]:
                continue
            if q[i] < self.lower[i] or self.upper[i] < q[i]:
                return False, ValidationReport(2, i)
        return True, None


class _PathValidator:
    method = 0

    def __init__(self, workspace, param, check_bounds=False):
        self.ws = workspace
        self.param = float(param)
        self.check_bounds = check_bounds
        self._ctor_check()

    def _ctor_check(self):
        pass

    def validate(self, path, reverse=False):
        """bool validate(const PathPtr_t&, bool reverse, PathPtr_t& validPart, PathValidationReportPtr_t&)"""
        r = self.validate_batch(path.initial.reshape(1, -1), path.end.reshape(1, -1), np.array([path.length]), reverse)
        if r["valid"][0]:
            return True, path, None
        tl = float(r["t_last_valid"][0])
        part = path.extract(tl, path.length, self.ws) if reverse else path.extract(0.0, tl, self.ws)
        return False, part, ValidationReport(r["reason"][0], r["pair"][0], parameter=float(r["t_fail"][0]))

    def validate_config(self, q):
        """PathValidation::validate(ConfigurationIn_t, report): zero-length StraightPath
        (reference src/path-validation.cc:36-49)."""
        ok, _, rep = self.validate(StraightPath(q, q, 0.0))
        return ok, rep

    def validate_batch(self, q1, q2, L, reverse=False):
        return self.ws.validate_edges(self.method, self.param, q1, q2, L, reverse=reverse,
                                      check_bounds=self.check_bounds)


class Discretized(_PathValidator):
    method = 0

    @classmethod
    def create(cls, workspace, stepSize, check_bounds=False):
        return cls(workspace, stepSize, check_bounds)


class Progressive(_PathValidator):
    method = 1

    def _ctor_check(self):  # progressive.cc:138-145
        if self.param <= 0:
            raise RuntimeError("tolerance should be positive for progressive continuous validation.")

    @classmethod
    def create(cls, workspace, tolerance):
        return cls(workspace, tolerance)


class Dichotomy(_PathValidator):
    method = 2

    def _ctor_check(self):  # dichotomy.cc:139-148
        if self.param != 0:
            raise RuntimeError("Dichotomy path validation method does not support penetration.")

    @classmethod
    def create(cls, workspace, tolerance=0.0):
        return cls(workspace, tolerance)
