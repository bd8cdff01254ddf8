"""Host statement of the optimiser with the oracle as evaluate (CPU only)."""
import numpy as np

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.registration import (RegistrationConfig, host_optimizer, params_to_index_matrix,
                                                    probe_params)
from oracle import native


def small_pair(shape=(20, 32, 40), gt=(2.0, -1.5, 2.5, 2.0, -1.5, 1.0)):
    sp = (2.0, 2.0, 2.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    center = np.array([0.0, 0.0, 0.0])
    Mgt = G.params_to_matrix(gt, center)
    fixed = make_phantom(shape, "CT", seed=4, spacing=sp, noise=False)
    moving = make_phantom(shape, "CT", seed=4, spacing=sp, noise=False, world_matrix=np.linalg.inv(Mgt))
    return g, center, fixed, moving


def oracle_evaluate(g, center, fixed, moving):
    p2w, w2p = G.index_to_physical(g), G.physical_to_index(g)

    def evaluate(plist):
        out = []
        for p in plist:
            s, n = native.reslice_ssd(fixed, moving, params_to_index_matrix(p, center, p2w, w2p), "itk")
            out.append(s / n if n else 1e300)
        return out
    return evaluate


def test_probe_set_layout():
    ps = probe_params(np.zeros(6), 0.5, np.array([1, 1, 1, 2, 2, 2.0]))
    assert len(ps) == 13 and np.all(ps[0] == 0)
    assert ps[1][0] == 0.5 and ps[2][0] == -0.5 and ps[7][3] == 1.0 and ps[12][5] == -1.0


def test_host_optimizer_reduces_cost_and_approaches_ground_truth():
    g, center, fixed, moving = small_pair()
    cfg = RegistrationConfig(metric="ssd", iterations=40, rot_step_deg=0.5, trans_step_mm=0.5)
    res = host_optimizer(oracle_evaluate(g, center, fixed, moving), np.zeros(6), cfg)
    assert res.iterations == 40 and res.evaluations == 13 * 40
    assert res.cost_history[-1] < 0.35 * res.cost_history[0]
    # accepted centres never get worse
    acc = [c for c, s0, s1 in zip(res.cost_history[1:], res.scale_history[:-1], res.scale_history[1:]) if s1 == s0]
    gt = np.array([2.0, -1.5, 2.5, 2.0, -1.5, 1.0])
    assert np.linalg.norm(res.params - gt) < np.linalg.norm(gt) * 0.5


def test_host_optimizer_stays_at_the_optimum():
    g, center, fixed, _ = small_pair()
    cfg = RegistrationConfig(metric="ssd", iterations=6)
    res = host_optimizer(oracle_evaluate(g, center, fixed, fixed), np.zeros(6), cfg)
    assert res.cost_history[0] == 0.0
    assert np.allclose(res.params, 0.0, atol=0.6)


def test_result_checkpoint_round_trip_and_resume(tmp_path):
    """the optimiser state persists as a small .npz; a resumed run starts at the saved parameters with the
    step lengths the first run had reached."""
    import numpy as np
    from medicalimageanalysis_b200.registration import RegistrationConfig, host_optimizer, resume_from

    def evaluate(probes):                          # a convex bowl with its minimum at (1, -2, 0.5, 3, 0, -1)
        t = np.array([1.0, -2.0, 0.5, 3.0, 0.0, -1.0])
        return [float(np.sum((np.asarray(p) - t) ** 2)) for p in probes]
    cfg = RegistrationConfig(iterations=25)
    first = host_optimizer(evaluate, np.zeros(6), cfg)
    first.save(tmp_path / "state")
    p0, cfg2 = resume_from(tmp_path / "state", cfg)
    assert np.array_equal(p0, first.params)
    s = float(first.scale_history[-1])
    assert cfg2.rot_step_deg == cfg.rot_step_deg * s and cfg2.min_step_scale * s == cfg.min_step_scale
    second = host_optimizer(evaluate, p0, cfg2)
    assert second.cost_history[-1] <= first.cost_history[-1]
    z = np.load(str(tmp_path / "state") + ".npz")
    assert int(z["iterations"]) == 25 and z["param_history"].shape == (25, 6)
