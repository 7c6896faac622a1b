"""Regenerates tests/golden/ from the read-only reference checkout (run in the build
container only; the GPU box has no /root/reference).

What is extracted (data only, no reference source code):
  * fixtures/: the .bc boundary-condition files, problem JSONs and the material JSON
    that the reference's scripts feed to pyVoxelFEM (inputs of every parity case);
  * quadrature_goldens.json: exact monomial integrals + per-variable degrees parsed from
    VoxelFEM/tests/tp_quadrature_{1,2,3}var_test.inl (the reference's only self-checking
    unit test, test_tp_gauss_quadrature.cc:44-73, tolerance 5e-16);
  * oc_trajectories.json: "Compliance loss" per OC step and the
    "objective, constraint, lambda estimate" lines of logs/slurm/gt/*.log.
"""
import json
import os
import re
import shutil

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def copy_inputs():
    dst = os.path.join(HERE, "fixtures")
    for sub in ("bcs/2d", "bcs/3d", "problems/2d", "problems/3d"):
        os.makedirs(os.path.join(dst, sub), exist_ok=True)
        for f in os.listdir(os.path.join(REF, sub)):
            shutil.copy(os.path.join(REF, sub, f), os.path.join(dst, sub, f))
    os.makedirs(os.path.join(dst, "materials"), exist_ok=True)
    shutil.copy(os.path.join(REF, "VoxelFEM/examples/materials/B9Creator.material"), os.path.join(dst, "materials"))


def quadrature():
    out = {}
    for nv in (1, 2, 3):
        txt = open(os.path.join(REF, "VoxelFEM/tests/tp_quadrature_%dvar_test.inl" % nv)).read()
        m = re.search(r"integrals%dVars\s*=\s*\{(.*?)\};" % nv, txt, re.S)
        integrals = []
        for tok in m.group(1).split(","):
            a, b = tok.strip().split("/")
            integrals.append([float(a.rstrip(".")), float(b)])  # stored as exact num/den
        m = re.search(r"varDegrees%dVars\s*=\s*\{(.*?)\};" % nv, txt, re.S)
        degs = [[int(v) for v in g.split(",")] for g in re.findall(r"\{([0-9, ]+)\}", m.group(1))]
        assert len(degs) == len(integrals)
        out["%dvar" % nv] = {"integrals_num_den": integrals, "degrees": degs}
    json.dump(out, open(os.path.join(HERE, "quadrature_goldens.json"), "w"))


def trajectories(max_steps=60):
    out = {}
    for log in ("2dMbb300x100", "2dBridge250x125", "c1001", "b1000"):
        txt = open(os.path.join(REF, "logs/slurm/gt/%s.log" % log)).read()
        cfg = re.search(r"VoxelFEM problem configs: (\{.*\})", txt).group(1)
        mgl = int(re.search(r"Multigrid levels: (\d+)", txt).group(1))
        comp = [float(v) for v in re.findall(r"Total Steps: \d+, Runtime: [0-9.]+, Compliance loss ([0-9.]+)", txt)]
        oc = [[float(a), float(b), float(c)] for a, b, c in
              re.findall(r"objective, constraint, lambda estimate: (\S+)\s+(\S+)\s+(\S+)", txt)]
        # the stdout lines of run k precede the stderr block in the captured log; keep the run's own lines
        out[log] = {"config": cfg, "multigrid_levels": mgl, "compliance": comp[:max_steps],
                    "compliance_step_1499": comp[1499] if len(comp) > 1499 else None,
                    "oc_lines_head": oc[:max_steps], "n_steps_logged": len(comp)}
    json.dump(out, open(os.path.join(HERE, "oc_trajectories.json"), "w"), indent=1)


if __name__ == "__main__":
    copy_inputs()
    quadrature()
    trajectories()
    print("fixtures written to", HERE)
