"""Several cfg2-sized queries in ONE fused launch per iteration (n_queries > 1): the serial tail (ranking, update, barriers,
launch turnaround) is paid once per launch, so points/s rise with the number of queries a GPU serves at a time.
usage: python scripts/batched_queries.py"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED, Q_START, Q_END
out = []
for Q in (1, 2, 3, 4, 8):
    case = joint_case(n_samples=4096, check_points=64, n_points=10, do_update=True, n_queries=Q)
    rng = np.random.default_rng(1)
    case.start = Q_START + rng.normal(0, 0.05, (Q, 7)); case.end = Q_END + rng.normal(0, 0.05, (Q, 7))
    case.start[0], case.end[0] = Q_START, Q_END
    h = capi.Handle(case.model, case.cfg); st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    case.prepare(h); h.save_state(); h.set_result_mirror(False)
    for i in range(10): h.iterate(SEED, i, 1, restored=True)
    torch.cuda.synchronize(); best = 1e9
    for rep in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for i in range(200): h.iterate(SEED, 10 + i, 1, restored=True)
        b.record(st); torch.cuda.synchronize(); best = min(best, a.elapsed_time(b) * 1e3 / 200)
    d = h.dims
    out.append(dict(queries=Q, us_per_iteration=round(best, 2), us_per_query_iteration=round(best / Q, 2), points_per_s=round(Q * 4096 * 64 / best * 1e6), grid=d.grid, cand_per_cta=d.cand_per_cta,
                    n_success=[int(x) for x in h.results()["n_success"]]))
    print(json.dumps(out[-1]), flush=True)
    h.close()
