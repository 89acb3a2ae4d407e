"""CPU tests: the oracle (`oracle/torch_oracle.py`) against the golden vectors recorded from the unmodified
reference (`tests/golden/generate_golden.py`). This is what pins the oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import torch_oracle as to
import golden_io as gio


def scene_from_golden(g, wo=False):
    tag = "_wo" if wo else ""
    env = str(g["env"])
    n = g["ado0"].shape[0]
    if env == "kalman":
        v0 = sigma = w = torch.ones((1, n, int(g["P"])))
        noise = None
    else:
        v0, sigma, w = gio.t(g["v0" + tag])[None], gio.t(g["sigma" + tag])[None], gio.t(g["weights" + tag])[None]
        noise = gio.t(g["noise" + tag])[None]
    return to.Scene(env, str(g["ego_kind"]), float(g["dt"]), gio.t(g["ego0"])[None], gio.t(g["ego_goal"])[None],
                    gio.t(g["ado0"])[None], gio.t(g["goal"])[None], v0, sigma, w, noise)


def test_agent_known_answers():
    """Reference dynamics KATs (test/test_agents.py:176-194, :241-260) and batched unroll / dx_du vectors."""
    k = gio.load("agent_kats")
    for kind, x0, u, x1 in zip(k["kat_kind"], k["kat_x0"], k["kat_u"], k["kat_x1"]):
        traj = to.ego_unroll(gio.t(x0)[None], gio.t(u)[None, None], 1.0, str(kind))
        assert np.allclose(traj[0, 1].numpy(), x1)
    for kind in ("single", "double"):
        for t_hor in (1, 5, 20, 33):
            x0, u = gio.t(k[f"{kind}_{t_hor}_x0"]), gio.t(k[f"{kind}_{t_hor}_u"]).requires_grad_(True)
            traj = to.ego_unroll(x0[None], u[None], 0.4, kind)[0]
            gio.assert_parity(traj.detach().numpy(), k[f"{kind}_{t_hor}_traj"][:, :4], gio.RTOL_ROLLOUT, "unroll")
            jac = torch.stack([torch.autograd.grad(x, u, retain_graph=True)[0].flatten() for x in traj.flatten()])
            ref = k[f"{kind}_{t_hor}_dxdu"].reshape(t_hor + 1, 5, 2 * t_hor)[:, :4].reshape(-1, 2 * t_hor)
            gio.assert_parity(jac.numpy(), ref, gio.RTOL_ROLLOUT, "dx_du")


@pytest.mark.parametrize("name", gio.case_names())
def test_rollout(name):
    g = gio.load(name)
    scene = scene_from_golden(g)
    ego_traj = to.ego_unroll(scene.ego0, gio.t(g["controls"])[None], scene.dt, scene.ego_kind)
    gio.assert_parity(ego_traj[0].numpy(), g["ego_traj"][:, :4], gio.RTOL_ROLLOUT, "ego_traj")
    mus, sigmas = to.rollout(scene.env_kind, scene.ado0, scene.goal, scene.v0, scene.sigma, scene.weights,
                             scene.noise, ego_traj, scene.dt)
    gio.assert_parity(mus[0, ..., 0:2].numpy(), g["mus"][..., 0:2], gio.RTOL_ROLLOUT, "mu_p")
    gio.assert_parity(mus[0, ..., 2:4].numpy(), g["mus"][..., 2:4], gio.RTOL_ROLLOUT, "mu_v")
    gio.assert_parity(sigmas[0].numpy(), g["sigmas"], gio.RTOL_ROLLOUT, "sigmas")
    scene_wo = scene_from_golden(g, wo=True)
    mus_wo, sig_wo = to.rollout(scene.env_kind, scene_wo.ado0, scene_wo.goal, scene_wo.v0, scene_wo.sigma,
                                scene_wo.weights, scene_wo.noise, int(g["T"]), scene.dt)
    gio.assert_parity(mus_wo[0].numpy(), g["mus_wo"], gio.RTOL_ROLLOUT, "mus_wo")
    gio.assert_parity(sig_wo[0].numpy(), g["sigmas_wo"], gio.RTOL_ROLLOUT, "sigmas_wo")


MODULE_WEIGHT = {"goal": 0.5, "acc": 1.0, "vel": 1.0, "pos": 1.0, "prob": 1.0}


@pytest.mark.parametrize("name", gio.case_names())
def test_module_objective_gradient(name):
    g = gio.load(name)
    scene = scene_from_golden(g)
    controls = gio.t(g["controls"])[None]
    mus_wo = gio.t(g["mus_wo"])[None]
    n = g["ado0"].shape[0]
    for key in ("goal", "acc", "vel", "pos", "prob"):
        if f"obj_{key}" not in g:
            continue
        for suffix, mask in (("", None), ("_sub2", torch.arange(n)[None] < 2)):
            if f"obj_{key}{suffix}" not in g:
                continue
            res = to.evaluate(scene, controls, mus_wo, {key: MODULE_WEIGHT[key]}, ado_mask=mask)
            col = ("goal", "acc", "vel", "pos", "prob").index(key)
            gio.assert_parity(res["obj"][0, col].numpy(), g[f"obj_{key}{suffix}"], gio.RTOL_ROLLOUT, f"obj_{key}{suffix}")
            gio.assert_parity(res["grad"][0].flatten().numpy(), g[f"grad_{key}{suffix}"], gio.RTOL_GRAD,
                              f"grad_{key}{suffix}")
            if suffix == "" and key != "goal" and not (key == "prob" and abs(float(g["raw_prob"])) > 120.0):
                gio.assert_parity(res["raw"][0, col].numpy(), g[f"raw_{key}"], gio.RTOL_ROLLOUT, f"raw_{key}")


@pytest.mark.parametrize("name", gio.case_names())
def test_constraints(name):
    g = gio.load(name)
    scene = scene_from_golden(g)
    t_hor = int(g["T"])
    ego_traj = to.ego_unroll(scene.ego0, gio.t(g["controls"])[None], scene.dt, scene.ego_kind)
    con = to.control_limit_raw(ego_traj, scene.dt, scene.ego_kind)[0].numpy() / to.ROBOT_ACC_MAX
    gio.assert_parity(con, g["con_control_limits"], gio.RTOL_ROLLOUT, "control limits")
    vio = to.violation(con, None, 1.0)
    gio.assert_parity(vio, g["vio_control_limits"], 1e-4, "control limit violation")
    spd = to.speed_limit_raw(ego_traj)[0].numpy() / to.ROBOT_SPEED_MAX
    gio.assert_parity(spd, g["con_speed_limits"], gio.RTOL_ROLLOUT, "speed limits")
    gio.assert_parity(to.violation(spd, -1.0, 1.0), g["vio_speed_limits"], 1e-4, "speed violation")
    # analytic Jacobians of the reference are constants of the dynamics: identity and the velocity rows of dx_du.
    if scene.ego_kind == "double":
        assert np.array_equal(g["jac_control_limits"].reshape(2 * t_hor, 2 * t_hor), np.eye(2 * t_hor))
    dxdu = g["dx_du"].reshape(t_hor + 1, 5, 2 * t_hor)
    gio.assert_parity(g["jac_speed_limits"].reshape(t_hor + 1, 2, 2 * t_hor), dxdu[:, 2:4], 1e-6, "speed jacobian")


def test_min_distance_jacobian():
    g = gio.load("c3_sf_n10_p4_t20")
    scene = scene_from_golden(g)
    con, jac = to.min_distance_jacobian(scene, gio.t(g["controls"])[None])
    gio.assert_parity(con.numpy(), g["con_min_distance"], gio.RTOL_ROLLOUT, "min distance")
    gio.assert_parity(jac.numpy().flatten(), g["jac_min_distance"], gio.RTOL_GRAD, "min distance jacobian")


def test_kalman_properties():
    """test/test_environment.py:254-269: constant mean velocity, non-decreasing variance."""
    g = gio.load("kalman_n5_t15")
    scene = scene_from_golden(g)
    mus, sigmas = to.rollout("kalman", scene.ado0, scene.goal, scene.v0, scene.sigma, scene.weights, None, 15, 0.4)
    assert torch.allclose(mus[0, :, :, 2:4], scene.ado0[0, :, None, 2:4].expand(-1, 15, -1))
    assert torch.all(sigmas[0, :, 1:] >= sigmas[0, :, :-1])
    pos_sig = to.kalman_pos_sigma(1, 5, 15)
    assert math.isclose(float(pos_sig[0, 0, 3, 0]), 1e-6 + 0.6, rel_tol=1e-5)


@pytest.mark.parametrize("name", gio.case_names())
def test_ellipsoid(name):
    """InteractionEllipsoidModule (modules/baselines/ellipsoid_interact.py): rows, violation, bounds and Jacobian."""
    g = gio.load(name)
    scene = scene_from_golden(g)
    con, jac = to.ellipsoid_jacobian(scene, gio.t(g["controls"])[None])
    gio.assert_parity(con.numpy(), g["con_ellipsoid"], gio.RTOL_ROLLOUT, "ellipsoid rows")
    if scene.env_kind == "kalman":   # not differentiable w.r.t. the ego: the reference returns zeros (optimization_module.py:298-301)
        assert not np.any(g["jac_ellipsoid"]) and g["jac_ellipsoid"].size == jac.numel()
    else:
        gio.assert_parity(jac.numpy().flatten(), g["jac_ellipsoid"], gio.RTOL_GRAD, "ellipsoid jacobian")
    gio.assert_parity(to.violation(con.numpy()[None], 1.0, None)[0], g["vio_ellipsoid"], 1e-4, "ellipsoid violation", atol=1e-6)
    assert np.all(g["lb_ellipsoid"] == 1.0) and np.all(np.isnan(g["ub_ellipsoid"]))


def test_attention_filters():
    """EuclideanModule / ClosestModule (mantrap/attention) on the recorded scenes, incl. the `~argmin` quirk and ties."""
    g = gio.load("scene_ops")
    quirk = 0
    for i in range(len(g["att_n"])):
        n = int(g["att_n"][i])
        ego, ado = gio.t(g["att_ego"][i])[None], gio.t(g["att_ado"][i][:n])[None]
        for kind in ("euclidean", "closest"):
            assert np.array_equal(to.attention_mask(kind, ego, ado)[0].numpy(), g[f"att_{kind}"][i][:n]), (i, kind)
        quirk += int(g["att_closest"][i].sum() > 1)
    assert quirk > 0          # the closest filter keeps more than one ado in some scenes: the reference's behaviour


def test_env_step():
    """GraphBasedEnvironment.step, five consecutive steps per recorded scene, with the recorded standard-normal draws."""
    g = gio.load("scene_ops")
    for c in range(len(g["step_n"])):
        n, ego_kind = int(g["step_n"][c]), str(g["step_ego_kind"][c])
        ego, ado = gio.t(g["step_ego0"][c])[None], gio.t(g["step_ado0"][c][:n])[None]
        for k in range(g["step_actions"].shape[1]):
            ego, ado = to.env_step(ego_kind, 0.4, ego, ado, gio.t(g["step_actions"][c][k])[None], gio.t(g["step_eps"][c][k][:n])[None])
            gio.assert_parity(ego[0].numpy(), g["step_ego_states"][c][k], gio.RTOL_ROLLOUT, f"ego {c}/{k}")
            gio.assert_parity(ado[0].numpy(), g["step_ado_states"][c][k][:n], gio.RTOL_ROLLOUT, f"ados {c}/{k}")


def test_scene_encoding_and_warm_start():
    """TrajOptSolver.encode and warm_start("encoding") (trajopt.py:466-499, :308-331) on the 40 recorded scenes."""
    g = gio.load("scene_ops")
    enc_db, sol_db = gio.t(g["ws_encoding_db"]), gio.t(g["ws_solution_db"])
    assert enc_db.shape == (100, 4) and sol_db.shape == (100, 10, 2)
    for i in range(len(g["ws_n"])):
        n, t_hor = int(g["ws_n"][i]), int(g["ws_T"][i])
        enc = to.scene_encoding(gio.t(g["ws_ego"][i])[None], gio.t(g["ws_goal"][i])[None], gio.t(g["ws_ado"][i][:n])[None])
        gio.assert_parity(enc[0].numpy(), g["ws_encoding"][i], gio.RTOL_ROLLOUT, f"encoding {i}", atol=1e-6)
        match, warm = to.warm_start_encoding(enc, enc_db, sol_db, t_hor)
        assert int(match[0]) == int(g["ws_match"][i])
        assert np.array_equal(warm[0].numpy(), g["ws_warm"][i][:t_hor])
