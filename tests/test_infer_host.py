"""Host-side pieces of the driver mirror (infer.py) and of the model classes that need no GPU."""
import numpy as np
import pytest

from scalable_stellar_inference_b200 import infer


def test_results_writer_roundtrip(tmp_path):
    """save_inference_data (scripts/model_star.py:191-196 writes arviz groups with to_netcdf): NetCDF-3 files per group."""
    from scipy.io import netcdf_file
    rng = np.random.default_rng(0)
    post = {"log_mass": rng.normal(size=(4, 50)), "y": rng.normal(size=(4, 50, 7)), "_unconstrained": rng.normal(size=(4, 50, 8))}
    stats = {"diverging": rng.random((4, 50)) < 0.1, "tree_depth": rng.integers(1, 10, (4, 50))}
    const = infer.flatten_dict({"M_H": {"loc": 0.1, "scale": 0.1}, "delta": {"cov": np.eye(4)}, "Av": 0.0})
    assert set(const) == {"M_H_loc", "M_H_scale", "delta_cov", "Av"}
    files = infer.save_inference_data(str(tmp_path / "out"), posterior=post, sample_stats=stats, constant_data=const,
                                      attrs=dict(star=10, inference_library="scalable_stellar_inference_b200"))
    assert sorted(f.rsplit("/", 1)[1] for f in files) == ["constant_data.nc", "posterior.nc", "sample_stats.nc"]
    with netcdf_file(str(tmp_path / "out" / "posterior.nc"), "r", mmap=False) as f:
        assert f.dimensions["chain"] == 4 and f.dimensions["draw"] == 50 and "_unconstrained" not in f.variables
        np.testing.assert_array_equal(f.variables["y"][:], post["y"])
        assert f.variables["y"].dimensions == ("chain", "draw", "y_dim_0")
        assert f.star == 10
    with netcdf_file(str(tmp_path / "out" / "sample_stats.nc"), "r", mmap=False) as f:
        np.testing.assert_array_equal(f.variables["diverging"][:], stats["diverging"].astype(np.int8))
    with netcdf_file(str(tmp_path / "out" / "constant_data.nc"), "r", mmap=False) as f:
        np.testing.assert_array_equal(f.variables["delta_cov"][:], np.eye(4))


def test_array_valued_prior_constants_are_rejected_not_averaged():
    """const['log_mass'] with per-star loc used to be collapsed with np.mean; only M_H may be per-star."""
    import scalable_stellar_inference_b200 as ssi
    with pytest.raises(ValueError, match="per-star prior parameters"):
        ssi.MultiStarModel(3, ["log_Teff"], const={"log_mass": dict(loc=np.array([0.0, 0.1, 0.2]), scale=0.3, low=-0.15, high=0.36)})


def test_nuts_mirror_rejects_what_is_not_implemented():
    with pytest.raises(NotImplementedError):
        infer.NUTS(object(), dense_mass=True)
