"""Device-resident automatic loss scaling (GPU twin of tests/test_lr_and_loss_scale.py): the skip
decision, the halving / doubling state machine and the gradient scale the update kernels use
all live in device memory; nothing is read back inside the step.
Reference: T/variable_mgr_util.py:51-139 (tf.cond inside the graph)."""
import types

import pytest
import torch

pytestmark = pytest.mark.gpu


def _engine(inc_every=2, scale=128.0, n=4096):
    from kf_benchmarks_b200 import params as params_lib
    from kf_benchmarks_b200.parallel.comm import SingleComm
    from kf_benchmarks_b200.parallel.strategy import UpdateEngine
    p = params_lib.make_params(model="trivial", use_fp16=True, fp16_enable_auto_loss_scale=True,
                               fp16_loss_scale=scale, fp16_inc_loss_scale_every_n=inc_every,
                               optimizer="sgd", variable_update="independent", weight_decay=0.0)
    dev = torch.device("cuda", 0)
    store = types.SimpleNamespace(
        device=dev, num_trainable_padded=n, w32=torch.ones(n, device=dev),
        grad=torch.zeros(n, device=dev), w16=torch.ones(n, device=dev, dtype=torch.float16),
        opt_slot=lambda name: torch.zeros(n, device=dev))
    model = types.SimpleNamespace(get_fp16_loss_scale=lambda: scale)
    return UpdateEngine(p, store, SingleComm(), model), store


def test_state_machine_matches_the_host_mirror_without_host_reads():
    from kf_benchmarks_b200.parallel.strategy import next_loss_scale
    eng, st = _engine(inc_every=2, scale=128.0)
    assert eng.device_ls and eng.ls_state is not None
    host_scale, host_steps, host_skipped = 128.0, 0, 0
    w_ref = torch.ones(st.num_trainable_padded)
    lr = 0.5
    # step pattern: clean, clean, clean (doubles), inf (skip + halve), clean, nan (skip), clean
    pattern = ["ok", "ok", "ok", "inf", "ok", "nan", "ok", "ok", "ok"]
    for i, kind in enumerate(pattern):
        eng.begin_step(lr)
        # gradients of (loss * scale): 1.0 * the scale the device currently holds
        kw = eng.loss_scale_kwargs()
        assert kw["loss_scale"] == 1.0 and kw["dyn_scale"].is_cuda
        st.grad.copy_(kw["dyn_scale"].expand(st.num_trainable_padded))   # device -> device
        if kind == "inf":
            st.grad[17] = float("inf")
        elif kind == "nan":
            st.grad[3] = float("nan")
        eng.finish_step()
        skipped = kind != "ok"
        if not skipped:
            w_ref -= lr * 1.0                                  # unscaled gradient is exactly 1
        host_scale, host_steps = next_loss_scale(host_scale, host_steps, skipped, 2)
        host_skipped += int(skipped)
    torch.cuda.synchronize()
    assert eng.loss_scale == host_scale
    assert eng.loss_scale_normal_steps == host_steps
    assert eng.skipped_steps == host_skipped == 2
    assert torch.allclose(st.w32.cpu(), w_ref, atol=1e-5)
    assert torch.isfinite(st.w32).all()
    assert int(eng.skip_flag.item()) == 0                      # cleared by the update kernel


def test_restore_and_inspection_go_through_the_device_state():
    eng, _ = _engine(scale=64.0)
    eng.loss_scale = 1024.0
    eng.loss_scale_normal_steps = 7
    assert eng.loss_scale == 1024.0 and eng.loss_scale_normal_steps == 7
    assert float(eng.ls_state[0]) == 1024.0
    eng.begin_step(0.1)                                         # creates dyn
    torch.cuda.synchronize()
    assert abs(float(eng.dyn[1]) - 1.0 / 1024.0) < 1e-9
