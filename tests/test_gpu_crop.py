"""GPU parity: crop + resize + normalise vs the oracle (bit-exact) and the golden cv2 / PIL outputs."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def _crops(g):
    img, xyxy = g["img"], g["xyxy"]
    return img, xyxy, [img[y1:y2, x1:x2] for x1, y1, x2, y2 in xyxy]


def test_cv_linear_frame_and_crops(rb, golden):
    from reidb200 import feature_extractor as FE, preprocess as P

    g = golden("preprocess")
    img, xyxy, crops = _crops(g)
    want = O.preprocess_deepsort(crops)
    got_frame = P.preprocess_frame(img, xyxy).cpu().numpy()
    assert got_frame.shape == (len(crops), 3, 256, 128)
    assert np.array_equal(got_frame, want), np.abs(got_frame - want).max()  # bit-exact vs the restatement
    ex = object.__new__(FE.Extractor)
    ex.size, ex.norm = (128, 256), P.DEEPSORT_NORM
    got_list = ex._preprocess(crops).cpu().numpy()
    assert np.array_equal(got_list, want)
    # vs cv2 itself (reference output): within 1 ulp of the /255 image scaled by 1/std
    assert np.abs(got_frame[g["cv_full_idx"]] - g["cv_full"]).max() <= 4.8e-7
    assert np.abs(got_frame[:, :, ::3, ::3] - g["cv_sub"]).max() <= 4.8e-7


def test_pil_antialias(rb, golden):
    from reidb200 import feature_extractor as FE, preprocess as P

    g = golden("preprocess")
    img, xyxy, crops = _crops(g)
    u8 = P.resize_pil_u8_crops(crops).cpu().numpy()
    assert np.array_equal(u8, g["pil_u8"])  # bit-exact integer resample
    got = FE.ReIDPreprocess()._preprocess(crops).cpu().numpy()
    assert np.abs(got[0] - g["pil0"]).max() <= 1e-6
    want = O.preprocess_boxmot(crops)
    assert np.abs(got - want).max() <= 1e-6
    got2 = FE.ReIDPreprocess().preprocess_frame(img, xyxy).cpu().numpy()
    assert np.array_equal(got, got2)


def test_mot_shaped_frame(rb):
    """1080p frame, 200 boxes incl. clamped ones: device crops equal the oracle on host slices."""
    from reidb200 import preprocess as P, synth

    fr = synth.mot_frame(M=200, T=4, seed=3)
    H, W = fr["img"].shape[:2]
    xyxy = P.xywh_to_xyxy(fr["det_xywh"], W, H)
    keep = (xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])
    xyxy = xyxy[keep]
    got = P.preprocess_frame(fr["img"], xyxy).cpu().numpy()
    sel = [0, 7, len(xyxy) // 2, len(xyxy) - 1]
    want = O.preprocess_deepsort([fr["img"][xyxy[i, 1]:xyxy[i, 3], xyxy[i, 0]:xyxy[i, 2]] for i in sel])
    assert np.array_equal(got[sel], want)


def test_large_small_and_odd_crops(rb):
    """Tile footprints beyond the staging buffer (direct-tap path), 1-pixel crops, odd offsets (unaligned 32-bit staging
    loads at every byte phase), non-power-of-two std (IEEE division path), every byte value."""
    from reidb200 import ops, preprocess as P

    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (700, 901, 3), dtype=np.uint8)  # odd row stride: rows start at every alignment
    img[0, :256, 0] = np.arange(256)
    xyxy = np.array([[0, 0, 901, 700], [3, 5, 880, 650], [1, 0, 257, 1], [10, 10, 11, 400], [7, 9, 8, 10], [2, 3, 131, 259],
                     [5, 1, 69, 129], [0, 0, 256, 2], [333, 100, 600, 699]], np.int32)
    crops = [img[y1:y2, x1:x2] for x1, y1, x2, y2 in xyxy]
    got = P.preprocess_frame(img, xyxy).cpu().numpy()
    assert np.array_equal(got, O.preprocess_deepsort(crops))
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    off, geom = P._geometry_from_xyxy(xyxy, img.shape[1])
    dev = torch.device("cuda")
    got2 = ops.crop_resize_norm(torch.from_numpy(img).to(dev), torch.from_numpy(off).to(dev), torch.from_numpy(geom).to(dev),
                                mean=mean, std=std).cpu().numpy()
    assert np.array_equal(got2, O.preprocess_deepsort(crops, mean=mean, std=std))


def test_empty_and_bad_crops(rb):
    from reidb200 import preprocess as P

    img = np.zeros((50, 60, 3), np.uint8)
    assert P.preprocess_frame(img, np.zeros((0, 4), np.int32)).shape == (0, 3, 256, 128)
    with pytest.raises(ValueError):
        P.preprocess_frame(img, np.array([[5, 5, 5, 20]], np.int32))
    with pytest.raises(ValueError):
        P.preprocess_crops([np.zeros((0, 4, 3), np.uint8)])


@pytest.mark.parametrize("kernel", ["v3", "v4", "v4p"])
@pytest.mark.parametrize("W", [420, 421, 423])
def test_cv_linear_kernel_variants_bit_equal(rb, monkeypatch, kernel, W):
    """The crop kernels (v3: thread = output column; v4: thread = four adjacent columns with 128-bit stores and the
    conversion-free byte -> float step; v4p, the default: v4 on packed f32x2 arithmetic with the source lines prefetched)
    against the oracle, bit for bit, on a frame that holds every byte value in
    every channel, crops that are up-scaled, down-scaled, one pixel wide, and clamped at the frame border, for frame widths
    whose row stride is and is not a multiple of four bytes."""
    from reidb200 import preprocess as P

    monkeypatch.delenv("REIDB200_CROP_V3", raising=False)
    monkeypatch.delenv("REIDB200_CROP_V4", raising=False)
    if kernel != "v4p":
        monkeypatch.setenv("REIDB200_CROP_V3" if kernel == "v3" else "REIDB200_CROP_V4", "1")
    rng = np.random.default_rng(11)
    H = 300
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    img[:256, 0, :] = np.arange(256, dtype=np.uint8)[:, None]   # every byte value in every channel
    img[0, :256, 1] = np.arange(256, dtype=np.uint8)
    boxes = np.array([[0, 0, 1, 256], [0, 0, 256, 1], [0, 0, W, 300], [17, 9, 145, 265], [100, 50, 131, 120], [3, 7, 303, 290],
                      [W - 20, 280, W, 300], [50, 60, 51, 61], [10, 10, 138, 266], [1, 0, 90, 300], [2, 1, 203, 299],
                      [W - 7, 0, W, 300]], np.int32)
    got = P.preprocess_frame(img, boxes).cpu().numpy()
    want = O.preprocess_deepsort([img[b[1]:b[3], b[0]:b[2]] for b in boxes])
    assert np.array_equal(got, want)
    got_i = P.preprocess_frame(img, boxes, norm=P.IMAGENET_NORM).cpu().numpy()   # std not a power of two: true division
    want_i = O.preprocess_deepsort([img[b[1]:b[3], b[0]:b[2]] for b in boxes], mean=P.IMAGENET_NORM[0], std=P.IMAGENET_NORM[1])
    assert np.array_equal(got_i, want_i)
