"""Golden for the free-running grad-sdf loop on NOISY data (BASELINE.json config C1: 5-sphere orbit with Kinect noise, 100
frames, 640x480, 2 cm voxels, trunc 10) — the regime in which the reference's own `|xi| < 1e-3` stopping rule is
borderline (SURVEY.md "three facts", 3).  Runs main_scan_3d.cpp:208-266 (frame 0 fused at identity; then optimize, update
iff converged) through
  * the reference's own build, OMP_THREAD_LIMIT=1 (its bit-reproducible mode): the golden flags / iteration counts / poses;
  * the reference's own build with its default OpenMP threading, RUNS times: its run-to-run spread on identical inputs;
  * the C port (oracle/gsdf_oracle.c), which must equal the first.
Needs /root/reference (oracle/_ref).  Writes tests/golden/c1_noisy_grad_loop.npz.
    python tests/golden/make_c1_loop.py [RUNS=3]
"""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
N, W, H, VS, TRUNC = 100, 640, 480, np.float32(0.02), np.float32(10)


def run_loop(kind):
    import importlib
    import gsdf_loader
    gsdf_loader.load()
    po = gsdf_loader.load_oracle()
    synth = importlib.import_module("gradient_sdf_b200.synth")
    frames, poses, K, _ = synth.make_sequence(N, W, H)
    scan = po.RefScan(0, VS, TRUNC, W, H, K) if kind == "ref" else po.Oracle(0, VS, TRUNC, W, H, K)
    ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
    scan.update(frames[0], ident)
    scan.set_pose(ident)
    conv, iters, traj = [1], [0], [ident.copy()]
    for i in range(1, N):
        c, it = scan.optimize(frames[i])
        if c:
            scan.update(frames[i], scan.get_pose())
        conv.append(int(c)); iters.append(int(it)); traj.append(np.array(scan.get_pose(), np.float32))
    return {"conv": conv, "iters": iters, "traj": np.stack(traj).tolist(), "voxels": int(scan.voxel_count())}


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        print(json.dumps(run_loop(sys.argv[2])))
        sys.exit(0)
    runs = int(sys.argv[1]) if len(sys.argv) > 1 else 3

    def child(kind, env_extra):
        env = dict(os.environ, **env_extra)
        out = subprocess.run([sys.executable, __file__, "--child", kind], env=env, capture_output=True, text=True, check=True).stdout
        return json.loads(out.strip().splitlines()[-1])

    one = child("ref", {"OMP_THREAD_LIMIT": "1"})
    port = child("port", {"OMP_THREAD_LIMIT": "1"})
    os.environ.pop("OMP_THREAD_LIMIT", None)
    multi = [child("ref", {}) for _ in range(runs)]  # the reference's default OpenMP threading
    np.savez_compressed(ROOT / "tests" / "golden" / "c1_noisy_grad_loop.npz",
                        conv=np.array(one["conv"], np.int8), iters=np.array(one["iters"], np.int16),
                        traj=np.array(one["traj"], np.float32), voxels=one["voxels"],
                        port_conv=np.array(port["conv"], np.int8), port_iters=np.array(port["iters"], np.int16),
                        port_traj=np.array(port["traj"], np.float32),
                        omp_conv=np.array([m["conv"] for m in multi], np.int8),
                        omp_iters=np.array([m["iters"] for m in multi], np.int16),
                        omp_traj=np.array([m["traj"] for m in multi], np.float32),
                        omp_voxels=np.array([m["voxels"] for m in multi]))
    print("1 thread: converged", sum(one["conv"]) - 1, "of", N - 1, "| port", sum(port["conv"]) - 1,
          "| default OMP runs", [sum(m["conv"]) - 1 for m in multi])
