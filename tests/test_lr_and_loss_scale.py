"""Learning-rate curves and the loss-scale state machine (T/benchmark_cnn.py:1037-1169,
T/models/resnet_model_legacy.py:297-328, T/variable_mgr_util.py:51-139; tests
T/benchmark_cnn_test.py:888-1002, T/models/resnet_model_test.py, T/variable_mgr_util_test.py)."""
import pytest

from kf_benchmarks_b200 import datasets, lr, params as P
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.parallel.strategy import next_loss_scale


def _resnet(**kw):
    p = P.make_params(model="resnet50", **kw)
    return p, model_config.get_model_config("resnet50", datasets.create_dataset(None, "imagenet"),
                                            p)


def test_resnet_default_schedule():
    p, m = _resnet()
    n = datasets.IMAGENET_NUM_TRAIN_IMAGES
    f = lambda step: lr.get_learning_rate(p, step, n, m, 256)  # noqa: E731
    assert f(0) == 0.0
    assert f(25021) == pytest.approx(0.128 * 25021 / 25022)
    assert f(25022) == pytest.approx(0.128)
    assert f(150136) == pytest.approx(0.128) and f(150137) == pytest.approx(0.0128)
    assert f(300273) == pytest.approx(0.0128) and f(300274) == pytest.approx(0.00128)
    assert f(400364) == pytest.approx(0.00128) and f(400365) == pytest.approx(0.000128)
    assert f(450410) == pytest.approx(0.000128) and f(450411) == pytest.approx(0.0000128)
    # linear scaling with the batch size
    assert lr.get_learning_rate(p, 50000, n, m, 512) == pytest.approx(0.256)


def test_resnet_replicated_divides_by_num_gpus():
    p, m = _resnet(variable_update="replicated", num_gpus=8)
    assert m.get_scaled_base_learning_rate(256) == pytest.approx(0.128 / 8)
    p, m = _resnet(variable_update="parameter_server", num_gpus=8)
    assert m.get_scaled_base_learning_rate(256) == pytest.approx(0.128)


def test_piecewise_warmup_decay_min():
    ds = datasets.create_dataset(None, "imagenet")
    m = model_config.get_model_config("trivial", ds, None)
    p = P.make_params(piecewise_learning_rate_schedule="0.5;2;0.25;4;0.125",
                      num_learning_rate_warmup_epochs=1)
    f = lambda s: lr.get_learning_rate(p, s, 1000, m, 100)  # 10 steps per epoch  # noqa: E731
    assert f(0) == 0.0 and f(5) == pytest.approx(0.25) and f(10) == 0.5
    assert f(19) == 0.5 and f(20) == 0.25 and f(39) == 0.25 and f(40) == 0.125
    p = P.make_params(init_learning_rate=1.0, num_epochs_per_decay=2, learning_rate_decay_factor=0.5,
                      minimum_learning_rate=0.2)
    g = lambda s: lr.get_learning_rate(p, s, 1000, m, 100)  # noqa: E731
    assert g(0) == 1.0 and g(19) == 1.0 and g(20) == 0.5 and g(40) == 0.25 and g(60) == 0.2
    with pytest.raises(ValueError):
        lr.get_learning_rate(P.make_params(piecewise_learning_rate_schedule="1;2"), 0, 1000, m, 100)
    with pytest.raises(ValueError):
        lr.get_learning_rate(P.make_params(piecewise_learning_rate_schedule="1;2;3",
                                           init_learning_rate=1.0), 0, 1000, m, 100)
    assert lr.get_learning_rate(P.make_params(), 7, 1000, m, 100) == 0.005  # model default


def test_num_batches_and_epochs():
    assert lr.get_num_batches_and_epochs(P.make_params(), 100, 1000) == (100, 10.0)
    assert lr.get_num_batches_and_epochs(P.make_params(num_batches=25), 100, 1000) == (25, 2.5)
    assert lr.get_num_batches_and_epochs(P.make_params(num_epochs=0.55), 100, 1000)[0] == 5
    with pytest.raises(ValueError):
        lr.get_num_batches_and_epochs(P.make_params(num_batches=1, num_epochs=1.0), 1, 1)


def test_loss_scale_state_machine():
    scale, steps = 128.0, 0
    for _ in range(3):                       # counts up to n clean steps ...
        scale, steps = next_loss_scale(scale, steps, False, 3)
    assert (scale, steps) == (128.0, 3)
    scale, steps = next_loss_scale(scale, steps, False, 3)   # ... then doubles
    assert (scale, steps) == (256.0, 0)
    scale, steps = next_loss_scale(scale, 2, True, 3)        # non-finite: halve, reset
    assert (scale, steps) == (128.0, 0)
