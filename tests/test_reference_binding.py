"""Route B of INTEGRATION.md as a test: the UNMODIFIED reference's solver frame driving the kernels.

`mantrap_b200.binding.bind_reference` subclasses the imported reference's environments (overriding only
`_compute_distributions`, mantrap/environment/base/particle.py:98-183) and interaction modules (overriding only
`compute_gradient_analytically` / `compute_jacobian_analytically`, mantrap/modules/base/optimization_module.py:181-200,
:334-353). On top of them this file runs the reference's OWN code, untouched:

  (a) `TrajOptSolver.objective / constraints` (mantrap/solver/base/trajopt.py:368-434) and `IPOPTSolver.gradient /
      jacobian` (mantrap/solver/ipopt.py:118-154) at the golden z under the golden seeds, compared with the fixtures the
      unmodified reference produced (tests/golden/*.npz);
  (b) the reference's own tests with the bound classes substituted: `test/test_solver.py::TestSolvers::test_log_query`
      (the 12 cases SURVEY §4 found runnable: RandomSearch / MonteCarloTreeSearch x {Kalman, PF, SF} x 2 attentions, a full
      closed-loop `solve()`), `test/test_modules.py::TestObjectives::test_gradient_analytical` (:84-110) and
      `TestConstraints::test_jacobian_analytical` (:262-287).

The reference cannot travel to the GPU box, so the whole file is skipped where /root/reference is absent; in the build
container it runs on the host emulation of the real kernel sources (tests/host_emul). The `cuda` parametrisation is the
same body on the CUDA library for a machine that has both a GPU and the reference.
"""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch

import golden_io as gio
import kernel_cases as kc
from oracle import reference_harness as rh

pytestmark = pytest.mark.skipif(not rh.reference_available(), reason="the reference is not mounted at /root/reference")

SCENE_CASES = [n for n in gio.case_names() if not n.startswith("kalman")]


@pytest.fixture(scope="module", params=["emul", pytest.param("cuda", marks=pytest.mark.gpu)])
def bound(request):
    from mantrap_b200 import _cabi, binding, engine, runtime
    mantrap = rh.import_reference()
    if request.param == "emul":
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul"))
        import build_emul
        eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build_emul.build())), "cpu")
    else:
        eng = engine.Engine("cuda:0")
    previous = runtime.set_default_engine(eng)
    ns = binding.bind_reference(mantrap, eng)
    ns.mantrap = mantrap
    yield ns
    runtime.set_default_engine(previous)


def build_bound_env(ns, g):
    """The fixture's scene in a BOUND reference environment (same construction as oracle/reference_harness.build_env)."""
    mantrap = ns.mantrap
    env_class = {"sf": ns.SocialForcesEnvironment, "pf": ns.PotentialFieldEnvironment}[str(g["env"])]
    ego_class = {"double": mantrap.agents.DoubleIntegratorDTAgent, "single": mantrap.agents.IntegratorDTAgent}[str(g["ego_kind"])]
    env = env_class(ego_position=gio.t(g["ego0"][0:2]), ego_velocity=gio.t(g["ego0"][2:4]), ego_type=ego_class, dt=float(g["dt"]))
    for i in range(g["ado0"].shape[0]):
        kwargs = {"goal": gio.t(g["goal"][i])} if str(g["env"]) == "sf" else {}
        env.add_ado(position=gio.t(g["ado0"][i, 0:2]), velocity=gio.t(g["ado0"][i, 2:4]), identifier=f"a{i:02d}", **kwargs)
    return env


def set_particles(ns, num_p):
    """The reference freezes P as a def-time default (SURVEY §5 "Config"); the goldens were made by patching that default."""
    for cls in (ns.SocialForcesEnvironment, ns.PotentialFieldEnvironment):
        cls._compute_distributions.__defaults__ = (num_p, True)


def solver_on_fixture_scene(ns, g, env, module_classes, seed_wo=None):
    """An unmodified `IPOPTSolver` whose internal environment is the fixture's scene. `TrajOptSolver.__init__` plans on
    `env.copy()`, and `GraphBasedEnvironment.copy` re-adds the ados WITHOUT their goals (graph_based.py:471-474 ->
    `SocialForcesEnvironment.add_ado(goal=zeros(2))`), so the copy is a different scene than the one the fixture's modules
    were called on: re-attach the goals and rebuild the modules on it (their ego-free baseline is drawn in `__init__`)."""
    mantrap = ns.mantrap
    t_hor = int(g["T"])
    solver = mantrap.solver.IPOPTSolver(env, goal=gio.t(g["ego_goal"]), t_planning=t_hor, modules=list(module_classes))
    for ado, src in zip(solver.env.ados, env.ados):
        ado.params.update(src.params)
    for cls in module_classes:
        if seed_wo is not None:
            torch.manual_seed(seed_wo)
        module = cls(env=solver.env, t_horizon=t_hor, goal=solver.goal)
        solver._module_dict[module.name] = module
    return solver


@pytest.mark.parametrize("name", SCENE_CASES)
def test_reference_solver_callbacks_on_golden(bound, name):
    """(a) objective / gradient of the unmodified solver frame at the golden z, module by module and summed."""
    ns, mantrap = bound, bound.mantrap
    g = gio.load(name)
    set_particles(ns, int(g["P"]))
    t_hor, n = int(g["T"]), g["ado0"].shape[0]
    z = np.asarray(g["controls"], dtype=np.float64).flatten()
    bound_modules = {"acc": ns.InteractionAccelerationModule, "vel": ns.InteractionVelocityModule,
                     "pos": ns.InteractionPositionModule, "prob": ns.InteractionProbabilityModule}
    env = build_bound_env(ns, g)
    for key, cls in bound_modules.items():
        if f"obj_{key}" not in g:
            continue
        solver = solver_on_fixture_scene(ns, g, env, [mantrap.modules.GoalNormModule, cls], seed_wo=int(g["seed_wo"]))
        f_obj, f_grad = kc.prob_floor(g, None) if key == "prob" else (0.0, 0.0)
        torch.manual_seed(int(g["seed"]))
        obj = solver.objective(z)                               # UNMODIFIED TrajOptSolver.objective
        expect = float(g["obj_goal"]) + float(g[f"obj_{key}"])
        gio.assert_parity(obj, expect, gio.RTOL_ROLLOUT, f"{name}: solver.objective goal+{key}", atol=gio.ATOL_OBJ + 4 * f_obj)
        torch.manual_seed(int(g["seed"]))
        grad = solver.gradient(z)                               # UNMODIFIED IPOPTSolver.gradient -> compute_gradient_analytically
        assert grad.shape == (2 * t_hor,)
        gio.assert_parity(grad, g["grad_goal"] + g[f"grad_{key}"], gio.RTOL_GRAD, f"{name}: solver.gradient goal+{key}",
                          atol=gio.ATOL + 4 * f_grad)
        # attention subset through the reference's ado_ids argument
        if f"obj_{key}_sub2" in g and n >= 3:
            sub = solver.env.ado_ids[:2]
            mask = (torch.arange(n)[None] < 2)
            f_obj, f_grad = kc.prob_floor(g, mask) if key == "prob" else (0.0, 0.0)
            torch.manual_seed(int(g["seed"]))
            obj = solver.objective(z, ado_ids=sub)
            gio.assert_parity(obj, float(g["obj_goal"]) + float(g[f"obj_{key}_sub2"]), gio.RTOL_ROLLOUT,
                              f"{name}: objective goal+{key} (2 ados)", atol=gio.ATOL_OBJ + 4 * f_obj)
            torch.manual_seed(int(g["seed"]))
            grad = solver.gradient(z, ado_ids=sub)
            gio.assert_parity(grad, g["grad_goal"] + g[f"grad_{key}_sub2"], gio.RTOL_GRAD,
                              f"{name}: gradient goal+{key} (2 ados)", atol=gio.ATOL + 4 * f_grad)


@pytest.mark.parametrize("name", SCENE_CASES)
def test_reference_solver_constraints_on_golden(bound, name):
    """(a) constraints / dense Jacobian of the unmodified solver frame: control | speed | ellipsoid rows."""
    ns, mantrap = bound, bound.mantrap
    g = gio.load(name)
    if str(g["ego_kind"]) == "single":
        pytest.skip("single-integrator control rows: the reference's own rows / Jacobian are inconsistent (DESIGN §5 (i))")
    set_particles(ns, int(g["P"]))
    t_hor = int(g["T"])
    z = np.asarray(g["controls"], dtype=np.float64).flatten()
    env = build_bound_env(ns, g)
    solver = solver_on_fixture_scene(ns, g, env, [mantrap.modules.GoalNormModule, mantrap.modules.ControlLimitModule,
                                                  mantrap.modules.SpeedLimitModule, ns.InteractionEllipsoidModule])
    torch.manual_seed(int(g["seed"]))
    cons = solver.constraints(z)                                # UNMODIFIED TrajOptSolver.constraints
    expect = np.concatenate([g["con_control_limits"], g["con_speed_limits"], g["con_ellipsoid"]])
    assert cons.shape == expect.shape
    n_fix = g["con_control_limits"].shape[0] + g["con_speed_limits"].shape[0]
    gio.assert_parity(cons[:n_fix], expect[:n_fix], gio.RTOL_ROLLOUT, f"{name}: control | speed rows")
    gio.assert_parity(cons[n_fix:], expect[n_fix:], 1e-4, f"{name}: ellipsoid rows", atol=1e-5)
    torch.manual_seed(int(g["seed"]))
    jac = solver.jacobian(z)                                    # UNMODIFIED IPOPTSolver.jacobian -> compute_jacobian_analytically
    expect = np.concatenate([g["jac_control_limits"], g["jac_speed_limits"], g["jac_ellipsoid"]])
    assert jac.shape == expect.shape
    gio.assert_parity(jac, expect, gio.RTOL_GRAD, f"{name}: dense Jacobian", atol=1e-5)


def _reference_test_module(name):
    test_dir = os.path.join(rh.REFERENCE_ROOT, "test")
    if test_dir not in sys.path:
        sys.path.insert(0, test_dir)
    return __import__(name)


@pytest.mark.parametrize("solver_name", ["RandomSearch", "MonteCarloTreeSearch"])
@pytest.mark.parametrize("env_name", ["KalmanEnvironment", "PotentialFieldEnvironment", "SocialForcesEnvironment"])
@pytest.mark.parametrize("attention_name", ["ClosestModule", "EuclideanModule"])
def test_reference_test_log_query(bound, solver_name, env_name, attention_name):
    """(b) test/test_solver.py:197-223 (`-k "test_log_query and not IPOPT and not Trajectron"`, 12 cases): the reference's
    own closed-loop `solve()` -- hard warm start, search iterations, `env.step` -- over the bound environments."""
    ns, mantrap = bound, bound.mantrap
    ref_tests = _reference_test_module("test_solver")
    set_particles(ns, mantrap.constants.ENV_NUM_PARTICLES)
    env_class = getattr(ns, env_name, None) or getattr(mantrap.environment, env_name)      # Kalman: the reference's own
    torch.manual_seed(0)
    ref_tests.TestSolvers.test_log_query(getattr(mantrap.solver.baselines, solver_name), env_class,
                                         getattr(mantrap.attention, attention_name), mantrap.constants.WARM_START_HARD)


def _seeded(cls, seed):
    """Test-local subclass pinning the particle draw of both sides of the reference's analytic-vs-autograd comparison (the
    reference's test compares two calls, each of which would otherwise draw its own particles)."""
    class Seeded(cls):
        def compute_gradient_analytically(self, *args, **kwargs):
            torch.manual_seed(seed)
            return super().compute_gradient_analytically(*args, **kwargs)

        def compute_jacobian_analytically(self, *args, **kwargs):
            torch.manual_seed(seed)
            return super().compute_jacobian_analytically(*args, **kwargs)

        def compute_objective(self, *args, **kwargs):
            torch.manual_seed(seed)
            return super().compute_objective(*args, **kwargs)

        def compute_constraint(self, *args, **kwargs):
            torch.manual_seed(seed)
            return super().compute_constraint(*args, **kwargs)
    Seeded.__name__ = cls.__name__
    return Seeded


@pytest.mark.parametrize("module_name", ["InteractionProbabilityModule", "InteractionPositionModule",
                                         "InteractionVelocityModule", "InteractionAccelerationModule"])
@pytest.mark.parametrize("env_name", ["PotentialFieldEnvironment", "SocialForcesEnvironment"])
def test_reference_test_gradient_analytical(bound, module_name, env_name):
    """(b) test/test_modules.py:84-110: the fused-kernel gradient (`compute_gradient_analytically`) against the reference's
    `compute_gradient_auto_grad` through the bound environment's autograd node, the reference's tolerance (atol 1e-2)."""
    ns, mantrap = bound, bound.mantrap
    ref_tests = _reference_test_module("test_modules")
    set_particles(ns, mantrap.constants.ENV_NUM_PARTICLES)
    torch.manual_seed(3)
    ref_tests.TestObjectives.test_gradient_analytical(_seeded(getattr(ns, module_name), 77), getattr(ns, env_name))


@pytest.mark.parametrize("env_name", ["PotentialFieldEnvironment", "SocialForcesEnvironment"])
def test_reference_test_jacobian_analytical(bound, env_name):
    """(b) test/test_modules.py:262-287 with the ellipsoid constraint: `mtb_constraint_jacobian` against the reference's
    row-by-row autograd Jacobian."""
    ns, mantrap = bound, bound.mantrap
    ref_tests = _reference_test_module("test_modules")
    set_particles(ns, mantrap.constants.ENV_NUM_PARTICLES)
    torch.manual_seed(5)
    ref_tests.TestConstraints.test_jacobian_analytical(_seeded(ns.InteractionEllipsoidModule, 78), getattr(ns, env_name))
