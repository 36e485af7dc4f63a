"""Command-line entry: ``python -m kf_benchmarks_b200.cli --model=resnet50 ...``

Same contract as the reference's ``tf_cnn_benchmarks.py`` (T/tf_cnn_benchmarks.py:38-64):
parse flags, reject positional arguments, ``setup``, build ``BenchmarkCNN``, print the
header, run. With ``--num_gpus=N`` (N > 1) and no rank environment, the command re-launches
itself as N ranks (one per GPU), the role ``kungfu-run -np N`` / ``mpirun`` play for the
reference; under ``torchrun`` it uses the ranks it was given. A barrier at exit mirrors
``kungfu.run_barrier()``.
"""
from __future__ import annotations

import os
import sys

from . import benchmark, flags, params as params_lib
from .utils import log_util


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else list(argv)
    if any(a in ("-h", "--help", "--helpfull") for a in argv):
        print(flags.help_text())
        return 0
    try:
        p = params_lib.make_params_from_flags(argv)
    except flags.FlagError as e:
        print("FATAL Flags parsing error: %s" % e, file=sys.stderr)
        return 2
    world = int(os.environ.get("WORLD_SIZE", "0"))
    want = p.num_gpus if p.variable_update not in ("kungfu", "horovod") else 0
    if world == 0 and want > 1:
        from .parallel import launcher
        return launcher.launch_local(["-m", "kf_benchmarks_b200.cli"] + argv, want)
    p = benchmark.setup(p)
    bench = benchmark.BenchmarkCNN(p)
    if bench.rank == 0:
        import torch
        log_util.log_fn("kf-benchmarks-b200 (PyTorch %s)" % torch.__version__)
    bench.print_info()
    bench.run()
    return 0


if __name__ == "__main__":
    sys.exit(main())
