"""Input staging against the reference's own host code path: cv2.resize (OpenCV is what
PyraPose/utils/image.py:308-323 calls) on the float32 preprocessed image, then the zero-padded batch of
PyraPose/preprocessing/generator.py:320-336.

Tolerance: OpenCV's float32 INTER_LINEAR kernels and the CUDA kernel evaluate the same expression
(p00*a0 + p01*a1)*b0 + (p10*a0 + p11*a1)*b1 with float coefficients; OpenCV's SIMD code may fuse multiply-adds, so
results may differ in the last bits: |diff| <= 3.1e-5 on pixel values of magnitude <= 152 (2 ulp = 2 x 1.53e-5), most pixels identical.
Sizes, scales and the padding are exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

cv2 = pytest.importorskip("cv2")


def _ref_pipeline(raw_images, min_side=480, max_side=640):
    from pyrapose_b200.utils import image as ui
    out, scales = [], []
    for raw in raw_images:
        x = ui.preprocess_image(raw)                                    # image.py:35-62
        s = ui.compute_resize_scale(x.shape, min_side, max_side)        # image.py:281-305
        out.append(cv2.resize(x, None, fx=s, fy=s))                     # image.py:320
        scales.append(s)
    return ui.compute_inputs(out), scales                               # generator.py:320-336


@pytest.mark.parametrize("shape", [(480, 640), (720, 1280), (600, 600), (1080, 1920), (375, 1242), (240, 320)])
def test_resize_image_matches_cv2(cuda_device, shape):
    from pyrapose_b200.utils import image as ui
    rng = np.random.default_rng(shape[0] + shape[1])
    raw = rng.integers(0, 256, shape + (3,)).astype(np.uint8)
    x = ui.preprocess_image(raw)
    got, scale = ui.resize_image(x)
    ref_scale = ui.compute_resize_scale(x.shape)
    ref = cv2.resize(x, None, fx=ref_scale, fy=ref_scale)
    assert scale == ref_scale and got.shape == ref.shape and got.dtype == np.float32
    d = np.abs(got - ref)
    print("resize %s -> %s: max |diff| %.2e, identical %.2f%%" % (shape, got.shape[:2], d.max(), 100 * (d == 0).mean()))
    assert d.max() <= 3.1e-5
    assert min(got.shape[:2]) <= 480 and max(got.shape[:2]) <= 640
    with pytest.raises(ValueError):
        ui.resize_image(raw)


def test_stage_batch_equals_host_pipeline(cuda_device):
    """A ragged group (the generator's batches hold images of different sizes): preprocess + resize + zero padding on
    the device from raw uint8 == the reference's three host steps."""
    from pyrapose_b200.utils import image as ui
    rng = np.random.default_rng(5)
    raws = [rng.integers(0, 256, s + (3,)).astype(np.uint8) for s in ((720, 1280), (480, 640), (600, 800), (500, 333))]
    ref, ref_scales = _ref_pipeline(raws)
    got, scales = ui.stage_batch(raws)
    got = got.cpu().numpy()
    assert scales == ref_scales and got.shape == ref.shape
    assert np.abs(got - ref).max() <= 3.1e-5
    # padding is exactly zero and the unscaled 480x640 image is bit-identical to preprocess_image
    assert got.shape[1:3] == (640, 640) and scales[3] == 1.28          # (500, 333) -> (640, 426)
    assert np.array_equal(got[3, :, 426:], np.zeros_like(got[3, :, 426:])) and np.abs(got[3, :, 425]).min() > 0
    assert np.array_equal(got[1, 480:], np.zeros_like(got[1, 480:]))
    assert np.array_equal(got[1, :480], ui.preprocess_image(raws[1]))


def test_predict_on_images_runs_the_staged_batch(cuda_device):
    from pyrapose_b200 import models
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils import image as ui
    C = 4
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict(spec.init_weights(C, seed=3, head_init="spread", randomize_bn=True))
    infer = models.convert_model(train)
    rng = np.random.default_rng(6)
    raws = [rng.integers(0, 256, s + (3,)).astype(np.uint8) for s in ((300, 400), (240, 320))]
    outs, scales = infer.predict_on_images(raws, min_side=120, max_side=160)
    batch, ref_scales = _ref_pipeline(raws, 120, 160)
    assert scales == ref_scales and batch.shape == (2, 120, 160, 3)
    staged, _ = ui.stage_batch(raws, 120, 160)
    want = infer.predict_on_batch(staged.cpu().numpy())
    for a, b in zip(outs, want):
        assert np.array_equal(a, b)
