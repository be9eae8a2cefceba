"""Per-opcode warp-instruction mix of every kernel in an ncu report (read without a GPU).
    python tools/ncu_opcodes.py gpurun_out/x.ncu-rep [top_n]
"""
import sys
sys.path.insert(0, "/opt/nvidia/nsight-compute/2025.2.1/extras/python")
import ncu_report

rep = ncu_report.load_report(sys.argv[1])
top = int(sys.argv[2]) if len(sys.argv) > 2 else 14
for r in range(rep.num_ranges()):
    rng = rep.range_by_idx(r)
    for a in range(rng.num_actions()):
        act = rng.action_by_idx(a)
        m = act.metric_by_name("sass__inst_executed_per_opcode")
        if m is None:
            continue
        tot = act.metric_by_name("smsp__inst_executed.sum").as_double()
        items = sorted(((m.as_uint64(i), m.correlation_ids().as_string(i)) for i in range(m.num_instances())), reverse=True)
        name = act.name()
        print(f"{name[:70]}  total {tot:.4g}")
        print("   " + "  ".join(f"{n}={100 * v / tot:.1f}%" for v, n in items[:top]))
