"""CPU: the numpy restatement of the reference's summaries (oracle/summary_oracle.py) on known answers."""
import numpy as np

import summary_oracle as SO


def test_known_answers():
    s = np.arange(1, 6, dtype=np.float32).reshape(1, 5, 1) * np.ones((1, 1, 2), np.float32)    # 1..5
    out = SO.summaries(s, truth=np.array([[3.0, 10.0]], np.float32), railed=np.array([1, 0, 0, 0, 1], np.uint8),
                       q=(0.0, 0.25, 0.5, 1.0))
    assert out["quantiles"][0, :, 0].tolist() == [1.0, 2.0, 3.0, 5.0]
    assert out["median"][0].tolist() == [3.0, 3.0] and out["lower_median"][0].tolist() == [3.0, 3.0]
    assert out["sbc_rank"][0].tolist() == [0.4, 1.0]
    assert out["rail_frac"].tolist() == [0.4]
    even = np.array([4, 1, 3, 2], np.float32).reshape(1, 4, 1)
    o2 = SO.summaries(even, q=(0.5,))
    assert o2["median"][0, 0] == 2.5 and o2["lower_median"][0, 0] == 2.0 and o2["quantiles"][0, 0, 0] == 2.5


def test_calib_block_keys():
    rng = np.random.default_rng(0)
    truth = np.abs(rng.standard_normal((64, 11))) + 1.0
    samples = (truth + 0.1 * rng.standard_normal((64, 11)))[:, None, :] + 0.1 * rng.standard_normal((64, 100, 11))
    blk = SO.calib_block(samples, truth)
    assert set(blk) == {"n", "cov50", "cov50_headline_mean", "cov90", "cov90_headline_mean", "mc_frac_err_median",
                        "dl_frac_err_median", "tc_abs_err_median_s", "sbc_ks_p"}
    assert 0.3 <= blk["cov50_headline_mean"] <= 0.7 and blk["cov90_headline_mean"] >= 0.8
