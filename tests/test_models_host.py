"""Host-side logic of the plugin surface (no GPU): data handling, parameters, constraints.

Mirrors the behaviours SURVEY.md Appendix A lists for the reference's BaseGPRModel /
GPflowGPRModel (GPSat/models/base_model.py:83-277, gpflow_models.py:339-494).
"""
import numpy as np
import pandas as pd
import pytest

from gpsat_b200 import _lib
from gpsat_b200.models import B200GPRModel, BaseGPRModel


def _df(n=20, seed=0):
    rng = np.random.default_rng(seed)
    return pd.DataFrame({"x": rng.uniform(-1e5, 1e5, n), "y": rng.uniform(-1e5, 1e5, n),
                         "t": rng.integers(-4, 5, n).astype(float), "obs": rng.normal(2.0, 1.0, n)})


def test_ctor_dataframe_scaling_and_local_mean():
    df = _df()
    raw = df.copy()
    m = B200GPRModel(data=df, obs_col="obs", coords_col=["x", "y", "t"], coords_scale=[5e4, 5e4, 1],
                     obs_mean="local", obs_scale=2.0)
    assert m.coords.shape == (20, 3) and m.obs.shape == (20, 1)
    np.testing.assert_allclose(m.coords, raw[["x", "y", "t"]].values / np.array([5e4, 5e4, 1]))
    np.testing.assert_allclose(m.obs[:, 0], (raw["obs"].values - raw["obs"].mean()) / 2.0)
    np.testing.assert_allclose(m.obs_mean, [raw["obs"].mean()])
    pd.testing.assert_frame_equal(df, raw)            # the caller's frame is not modified
    assert m.param_names == ["lengthscales", "kernel_variance", "likelihood_variance"]
    assert m.coords_col == ["x", "y", "t"] and m.obs_col == ["obs"]
    assert isinstance(m.cpu_name, str)


def test_obs_mean_other_than_local_is_zero():
    """base_model.py:195-200: numeric obs_mean values are discarded."""
    m = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y"], obs_mean=5.0)
    np.testing.assert_array_equal(m.obs_mean, [[0]])


def test_ctor_arrays_and_asserts():
    rng = np.random.default_rng(1)
    m = B200GPRModel(coords=rng.uniform(size=10), obs=rng.normal(size=10))
    assert m.coords.shape == (10, 1) and m.obs.shape == (10, 1) and m.coords_col == [0] and m.obs_col == [0]
    with pytest.raises(AssertionError):
        B200GPRModel(coords=rng.uniform(size=(10, 2)), obs=rng.normal(size=9))
    bad = rng.normal(size=10)
    bad[3] = np.nan
    with pytest.raises(AssertionError):
        B200GPRModel(coords=rng.uniform(size=(10, 2)), obs=bad)
    with pytest.raises(AssertionError):
        B200GPRModel(data=_df(), obs_col="obs")          # coords_col missing
    with pytest.raises(NotImplementedError):
        B200GPRModel(data=_df(), obs_col="obs", coords_col=["x"], kernel="Cosine")
    with pytest.raises(NotImplementedError):
        B200GPRModel(data=_df(), obs_col="obs", coords_col=["x"], mean_function="Constant")
    # junk keyword arguments are swallowed like the reference's **kwargs (IS2_GPSat_train.py:831-835)
    B200GPRModel(data=_df(), obs_col="obs", coords_col=["x"], jitter=1e-4, likelihood_variance=0.1)


def test_defaults_and_get_set():
    m = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y", "t"])
    p = m.get_parameters()
    np.testing.assert_array_equal(p["lengthscales"], np.ones(3))        # gpflow_models.py:129-131
    assert p["kernel_variance"] == 1.0 and p["likelihood_variance"] == 1.0
    assert isinstance(p["kernel_variance"], float) and isinstance(p["likelihood_variance"], float)
    m.set_parameters(lengthscales=[2.0, 3.0, 4.0], kernel_variance=np.array([0.5]), likelihood_variance=0.01)
    assert m.get_parameters("kernel_variance", return_dict=False) == [0.5]
    np.testing.assert_array_equal(m.get_lengthscales(), [2.0, 3.0, 4.0])
    m.set_lengthscales(1.5)
    np.testing.assert_array_equal(m.get_lengthscales(), [1.5] * 3)
    with pytest.raises(AssertionError):
        m.set_kernel_variance(np.array([1.0, 2.0]))
    with pytest.raises(AssertionError):
        m.get_parameters("nope")
    with pytest.raises(AssertionError):
        m.set_parameters(nope=1.0)
    with pytest.warns(UserWarning):
        m.set_likelihood_variance(1e-9)                                  # gpflow_models.py:404-411
    assert m.get_likelihood_variance() >= 1e-6
    with pytest.raises(ValueError):
        m.set_kernel_variance(-1.0)
    k = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y"],
                     kernel_kwargs={"lengthscales": [3.0, 4.0], "variance": 2.0}, noise_variance=0.3)
    assert k.get_kernel_variance() == 2.0 and k.get_likelihood_variance() == 0.3
    np.testing.assert_array_equal(k.get_lengthscales(), [3.0, 4.0])


def test_constraints_scale_and_move_within_tol():
    """gpflow_models.py:459-494 and the driver call local_experts.py:1110-1115."""
    m = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y", "t"], coords_scale=[5e4, 5e4, 1])
    m.set_lengthscales([1.0, 20.0, 1e-9])
    cons = {"lengthscales": {"low": [1e-8, 1e-8, 1e-8], "high": [6e5, 6e5, 9], "scale": True},
            "kernel_variance": {"low": 0.5, "high": 0.9},
            "likelihood_variance": {"low": 1e-4, "high": 0.1}}
    m.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
    ls = m.get_lengthscales()
    high = np.array([12.0, 12.0, 9.0])
    low = np.array([2e-13, 2e-13, 1e-8])
    np.testing.assert_allclose(ls, [1.0, 12.0 - 1e-2, 1e-8 + 1e-2])
    np.testing.assert_array_equal(m._tkind, [_lib.TRANSFORM_SIGMOID] * 5)
    np.testing.assert_allclose(m._ta[:3], low)
    np.testing.assert_allclose(m._tb[:3], high)
    assert m.get_kernel_variance() == pytest.approx(0.9 - 1e-2)          # 1.0 clamped to high - tol
    assert m.get_likelihood_variance() == pytest.approx(0.1 - 1e-2)
    # tol is capped at half the narrowest interval (gpflow_models.py:473-475)
    m2 = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x"])
    m2.set_kernel_variance_constraints(low=0.999, high=1.001, tol=1e-2)
    assert m2.get_kernel_variance() == pytest.approx(1.0)
    with pytest.raises(AssertionError):
        m2.set_lengthscales_constraints(low=[1e-3, 1e-3], high=[1.0, 1.0])     # wrong length
    with pytest.raises(AssertionError):
        m2.set_lengthscales_constraints(low=2.0, high=1.0)
    # values can no longer be assigned outside the sigmoid's interval
    with pytest.raises(ValueError):
        m2.set_kernel_variance(5.0)


def test_fixed_params_persist_and_bad_optimiser_args():
    m = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y"])
    m._fix_hyperparameters(["likelihood_variance", "bogus"])
    np.testing.assert_array_equal(m._trainable, [True, True, True, False])
    with pytest.raises(NotImplementedError):
        m.optimise_parameters(method="BFGS")
    with pytest.raises(TypeError):
        m.optimise_parameters(bogus_option=1)


def test_predict_input_forms():
    m = B200GPRModel(data=_df(), obs_col="obs", coords_col=["x", "y", "t"], coords_scale=[5e4, 5e4, 1])
    one = m._prep_pred_coords(np.array([1e5, 5e4, 2.0]))            # 1-d array = one point (gpflow_models.py:220-221)
    np.testing.assert_allclose(one, [[2.0, 1.0, 2.0]])
    fr = m._prep_pred_coords(pd.DataFrame({"t": [1.0], "y": [5e4], "x": [0.0], "extra": [7]}))
    np.testing.assert_allclose(fr, [[0.0, 1.0, 1.0]])              # columns picked by coords_col
    ls = m._prep_pred_coords([[5e4, 5e4, 1.0], [0, 0, 0]], apply_scale=False)
    np.testing.assert_allclose(ls, [[5e4, 5e4, 1.0], [0, 0, 0]])
    np.testing.assert_array_equal(m._f_bar(3), [0, 0, 0])


def test_base_class_requires_get_set():
    class Broken(BaseGPRModel):
        param_names = ["a"]
        def predict(self, coords): ...
        def optimise_parameters(self): ...
        def get_objective_function_value(self): ...
    with pytest.raises(AttributeError):
        Broken(coords=np.zeros((3, 1)), obs=np.zeros(3))


def test_sgpr_model_inducing_points():
    """gpflow_models.py:805-819: M random rows of the scaled coordinates (numpy's global generator),
    or all of them when there are fewer observations; get / set round trip; param_names."""
    from gpsat_b200.models import B200SGPRModel
    df = _df(40)
    np.random.seed(3)
    m = B200SGPRModel(data=df, obs_col="obs", coords_col=["x", "y", "t"], coords_scale=[5e4, 5e4, 1],
                      num_inducing_points=10)
    Z = m.get_inducing_points()
    assert Z.shape == (10, 3)
    rows = {tuple(r) for r in m.coords}
    assert all(tuple(z) in rows for z in Z)                          # a sub-sample of the scaled data
    np.random.seed(3)
    expect = m.coords.copy()
    np.random.shuffle(expect)
    np.testing.assert_array_equal(Z, expect[:10])                     # same draw as the reference's
    assert m.param_names == ["lengthscales", "kernel_variance", "likelihood_variance", "inducing_points"]
    m.set_inducing_points(Z[:4] + 0.5)
    np.testing.assert_array_equal(m.get_parameters("inducing_points")["inducing_points"], Z[:4] + 0.5)
    with pytest.raises(AssertionError):
        m.set_inducing_points(np.zeros((3, 2)))
    small = B200SGPRModel(data=_df(8), obs_col="obs", coords_col=["x", "y"])        # default M = 500 > n
    np.testing.assert_array_equal(small.get_inducing_points(), small.coords)
    with pytest.raises(AssertionError):
        B200SGPRModel(data=_df(8), obs_col="obs", coords_col=["x", "y"], num_inducing_points=None)


def test_vff_model_domain_and_parameters():
    """vff_model.py:174-258: domain box widened to hold the data, per-dimension kernels, product
    variance, D-th-root constraints, the reference's last-kernel quirk for fixed_params."""
    from gpsat_b200.models import B200VFFModel
    df = _df(30)
    m = B200VFFModel(data=df, obs_col="obs", coords_col=["x", "y"], coords_scale=[5e4, 5e4],
                     num_inducing_features=[6, 4], domain_size=[3e5, 1e3], expert_loc=np.array([0.0, 0.0]))
    assert m.param_names == ["lengthscales", "kernel_variance", "likelihood_variance"]
    np.testing.assert_array_equal(m.vff_nfreq, [6, 4])
    assert m.num_features == 11 * 7
    # dimension 0: the requested box (+-3e5 / 5e4 = +-6) contains the data; dimension 1: too small, widened
    assert m.vff_a[0] == -6.0 and m.vff_b[0] == 6.0
    assert m.vff_a[1] == m.coords[:, 1].min() - 1e-8 and m.vff_b[1] == m.coords[:, 1].max() + 1e-8
    assert np.all(m.coords > m.vff_a) and np.all(m.coords < m.vff_b)
    np.testing.assert_array_equal(m.get_lengthscales(), [1.0, 1.0])
    assert m.get_kernel_variance() == 1.0 and m.get_likelihood_variance() == 1.0
    m.set_kernel_variance(4.0)
    np.testing.assert_allclose(m._theta[2:4], [2.0, 2.0])
    assert abs(m.get_kernel_variance() - 4.0) < 1e-14
    m.set_lengthscales([0.5, 2.0])
    m.set_parameter_constraints({"lengthscales": {"low": [1e-3, 1e-3], "high": [1e5, 2e5], "scale": True},
                                 "kernel_variance": {"low": 1e-4, "high": 9.0}})
    np.testing.assert_allclose(m._tb[:2], [2.0, 4.0])                 # divided by coords_scale
    np.testing.assert_allclose(m._ta[2:4], [1e-2, 1e-2])              # low ** (1 / D)
    np.testing.assert_allclose(m._tb[2:4], [3.0, 3.0])
    m._fix_hyperparameters(["lengthscales", "kernel_variance"])
    np.testing.assert_array_equal(m._trainable, [True, False, True, False, True])
    one = B200VFFModel(data=df, obs_col="obs", coords_col=["x"], num_inducing_features=5)
    assert one.num_features == 9 and one.vff_a[0] == one.coords.min() - 1e-8        # domain_size=None
    # update_obs_data: the box is kept where the new data fits and widened where it does not
    a0, b0 = m.vff_a.copy(), m.vff_b.copy()
    df2 = df.copy()
    df2["y"] = df2["y"] * 3.0
    m.update_obs_data(data=df2, obs_col="obs", coords_col=["x", "y"], coords_scale=[5e4, 5e4])
    assert m.vff_a[0] == a0[0] and m.vff_b[0] == b0[0]
    assert np.all(m.coords > m.vff_a) and np.all(m.coords < m.vff_b)
    assert m.vff_a[1] <= a0[1] and m.vff_b[1] >= b0[1]
    with pytest.raises(NotImplementedError):
        B200VFFModel(data=df, obs_col="obs", coords_col=["x"], num_inducing_features=5, kernel="RBF")
    with pytest.raises(AssertionError):
        B200VFFModel(data=df, obs_col="obs", coords_col=["x"])
