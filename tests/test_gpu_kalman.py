"""GPU parity, N3: batched device-resident Kalman filter (csrc/kalman.cu) against the recorded sequence of
the reference's StrongSortKalmanFilterAdapter (tests/golden/kalman.npz) and the oracle at MOT20 size."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def test_sequence_golden(rb, golden):
    from reidb200.kalman import BatchedStrongSortKF

    g = golden("kalman")
    T = g["z0"].shape[0]
    slots = np.arange(T, dtype=np.int32) * 3 + 1  # scattered slots of a larger table
    kf = BatchedStrongSortKF(max_tracks=64)
    kf.initiate(slots, g["z0"])
    x0 = kf.mean[torch.as_tensor(slots).long()].cpu().numpy()
    assert np.array_equal(x0[:, :4], g["z0"]) and not x0[:, 4:].any()
    for f in range(g["matched"].shape[0]):
        kf.predict(slots)
        m = np.nonzero(g["matched"][f])[0]
        kf.update(slots[m], g["meas"][f, m].astype(np.float64), g["conf"][f, m])
        mean, SI = kf.gate_state(slots)
        P = kf.covariance[torch.as_tensor(slots).long()].cpu().numpy()
        np.testing.assert_allclose(mean.cpu().numpy(), g["x"][f], rtol=1e-10, atol=1e-8)
        np.testing.assert_allclose(P, g["P"][f], rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(SI.cpu().numpy().reshape(T, 4, 4), g["SI"][f], rtol=1e-8, atol=1e-12)
    untouched = np.setdiff1d(np.arange(64), slots)
    assert not kf.mean[torch.as_tensor(untouched).long()].any()


def test_predict_is_bit_exact(rb):
    """dt = 1: every product with F's 0/1 entries is exact in numpy's BLAS too."""
    from reidb200.kalman import BatchedStrongSortKF

    rng = np.random.default_rng(2)
    T = 300
    z = np.stack([rng.uniform(0, 1900, T), rng.uniform(0, 1000, T), rng.uniform(0.2, 0.7, T), rng.uniform(40, 400, T)], 1)
    kf = BatchedStrongSortKF(max_tracks=T)
    kf.initiate(np.arange(T), z)
    st = [O.kf_initiate(z[t]) for t in range(T)]
    for _ in range(3):
        kf.predict(np.arange(T))
        st = [O.kf_predict(x, P) + (SI,) for x, P, SI in st]
    assert np.array_equal(kf.mean.cpu().numpy(), np.stack([s[0] for s in st]))
    assert np.array_equal(kf.covariance.cpu().numpy(), np.stack([s[1] for s in st]))


def test_gate_state_feeds_association(rb):
    """mean / SI stay on the device: frame_costs with the filter's tensors equals frame_costs with host copies."""
    from reidb200 import matching, synth, tracker
    from reidb200.kalman import BatchedStrongSortKF

    fr = synth.mot_frame(M=12, T=20, budget=10, D=64, H=270, W=480, seed=1)
    T = len(fr["samples"])
    metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 10, max_tracks=T)
    metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), [len(s) for s in fr["samples"]]), list(range(T)))
    kf = BatchedStrongSortKF(max_tracks=T)
    kf.initiate(np.arange(T), fr["trk_mean"][:, :4])
    kf.predict(np.arange(T))
    kf.update(np.arange(T), fr["trk_mean"][:, :4] + 1.0, np.full(T, 0.5))
    kf.predict(np.arange(T))
    mean, SI = kf.gate_state(np.arange(T))
    a1, _, i1 = tracker.frame_costs(metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), np.ones(T, bool), mean, SI,
                                    fr["trk_tsu"], 0.995, 0.2, 0.7)
    a2, _, i2 = tracker.frame_costs(metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), np.ones(T, bool),
                                    mean.cpu().numpy(), SI.cpu().numpy(), fr["trk_tsu"], 0.995, 0.2, 0.7)
    assert torch.equal(a1, a2) and torch.equal(i1, i2)
