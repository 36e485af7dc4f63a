"""Learning-rate rules (T/benchmark_cnn.py:1037-1169): model default schedule, or
--init_learning_rate with staircase exponential decay and a floor, or
--piecewise_learning_rate_schedule, each with optional linear warm-up. Evaluated on the
host once per step; the value reaches the fused update kernel as a scalar."""
from __future__ import annotations

_DEFAULT_NUM_BATCHES = 100


def get_num_batches_and_epochs(params, batch_size, num_examples_per_epoch):
    if params.num_batches and params.num_epochs:
        raise ValueError("At most one of --num_batches and --num_epochs may be specified.")
    if params.num_epochs:
        num_batches = int(float(params.num_epochs) * num_examples_per_epoch / batch_size)
    else:
        num_batches = params.num_batches or _DEFAULT_NUM_BATCHES
    num_epochs = num_batches * batch_size / float(num_examples_per_epoch)
    return num_batches, num_epochs


def parse_piecewise_schedule(schedule: str, num_batches_per_epoch: float):
    pieces = schedule.split(";")
    if len(pieces) % 2 == 0:
        raise ValueError("--piecewise_learning_rate_schedule must have an odd number of "
                         "components")
    values, boundaries = [], []
    for i, piece in enumerate(pieces):
        if i % 2 == 0:
            try:
                values.append(float(piece))
            except ValueError:
                raise ValueError("Invalid learning rate: " + piece)
        else:
            try:
                boundaries.append(int(int(piece) * num_batches_per_epoch) - 1)
            except ValueError:
                raise ValueError("Invalid epoch: " + piece)
    return values, boundaries


def piecewise_constant(step: int, boundaries, values) -> float:
    for b, v in zip(boundaries, values):
        if step <= b:
            return v
    return values[-1]


def get_learning_rate(params, global_step: int, num_examples_per_epoch, model, batch_size) -> float:
    num_batches_per_epoch = float(num_examples_per_epoch) / batch_size
    if params.piecewise_learning_rate_schedule:
        if (params.init_learning_rate is not None or params.learning_rate_decay_factor or
                params.minimum_learning_rate or params.num_epochs_per_decay):
            raise ValueError("No other learning rate-related flags can be specified if "
                             "--piecewise_learning_rate_schedule is specified")
        values, boundaries = parse_piecewise_schedule(params.piecewise_learning_rate_schedule,
                                                      num_batches_per_epoch)
        lr = piecewise_constant(global_step, boundaries, values)
    elif params.init_learning_rate is not None:
        lr = params.init_learning_rate
        if params.num_epochs_per_decay > 0 and params.learning_rate_decay_factor > 0:
            decay_steps = int(num_batches_per_epoch * params.num_epochs_per_decay)
            lr = params.init_learning_rate * (
                params.learning_rate_decay_factor ** (global_step // max(decay_steps, 1)))
            if params.minimum_learning_rate != 0.0:
                lr = max(lr, params.minimum_learning_rate)
    else:
        lr = model.get_learning_rate(global_step, batch_size)
    if params.num_learning_rate_warmup_epochs > 0 and (
            params.init_learning_rate is not None or params.piecewise_learning_rate_schedule):
        warmup_steps = int(num_batches_per_epoch * params.num_learning_rate_warmup_epochs)
        init_lr = params.init_learning_rate
        if init_lr is None:
            init_lr = float(params.piecewise_learning_rate_schedule.split(";")[0])
        if global_step < warmup_steps:
            lr = init_lr * float(global_step) / float(warmup_steps)
    return float(lr)
