"""One rank of a multi-process CPU run (spawned by tests through the launcher): trains a few
steps with the given params and dumps this rank's final variables."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from kf_benchmarks_b200 import benchmark, params as P  # noqa: E402
from kf_benchmarks_b200.utils import log_util  # noqa: E402


def main():
    kw = json.loads(sys.argv[1])
    out_dir = sys.argv[2]
    lines = []
    log_util.log_fn = lambda s: lines.append(s)
    p = benchmark.setup(P.make_params(**kw))
    bench = benchmark.BenchmarkCNN(p)
    bench.print_info()
    stats = bench.run()
    rank = bench.rank
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), w=bench.store.w32.cpu().numpy(),
             state=bench.store.state.cpu().numpy(), loss=stats["last_average_loss"],
             ips=stats["images_per_sec"])
    with open(os.path.join(out_dir, "rank%d.log" % rank), "w") as f:
        f.write("\n".join(lines))


if __name__ == "__main__":
    main()
