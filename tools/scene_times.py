"""Wall time of the UNMODIFIED config scenes (tests/golden/scenes/*.py = the reference's files) through the three executables:
oracle/_ref/manta (the reference, all host threads), integration/_build/manta_b200 (reference executable, hot path bound to the
C-ABI) and python -m mantaflow_b200 (ctypes mirror).  Two step counts per scene give the per-step time without start-up cost.
    python tools/scene_times.py > profiles/<tag>_scene_times.json"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tests", "golden", "scene_driver.py")
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
EXES = {"reference": [os.path.join(ROOT, "oracle", "_ref", "manta")],
        "manta_b200": [os.path.join(ROOT, "integration", "_build", "manta_b200")],
        "python -m mantaflow_b200": [sys.executable, "-m", "mantaflow_b200"]}


def run(exe, scene, steps, cwd):
    env = dict(os.environ, MF_SCENE=os.path.join(SCENES, scene), MF_SCENE_STEPS=str(steps), MF_SCENE_OUT=cwd, MF_SCENE_SAVE="",
               MANTA_DISABLE_UI="1", OMP_NUM_THREADS=str(os.cpu_count() or 1), PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    t0 = time.perf_counter()
    p = subprocess.run(exe + [DRIVER], env=env, cwd=cwd, capture_output=True, text=True, timeout=3000)
    assert p.returncode == 0 and "Traceback" not in p.stderr, p.stdout[-1000:] + p.stderr[-1000:]
    return time.perf_counter() - t0


def main():
    out = {}
    for scene, (a, b) in {"simpleplume.py": (40, 240), "karman.py": (200, 1400), "flip01_simple.py": (200, 1400)}.items():
        out[scene] = {}
        with tempfile.TemporaryDirectory() as cwd:
            for name, exe in EXES.items():
                if not os.path.exists(exe[0]):
                    continue
                run(exe, scene, 2, cwd)                     # warm (noise tile file, page cache)
                ta, tb = run(exe, scene, a, cwd), run(exe, scene, b, cwd)
                out[scene][name] = {"steps": [a, b], "wall_s": [ta, tb], "ms_per_step": 1e3 * (tb - ta) / (b - a)}
                sys.stderr.write("%s %s %.2f ms/step\n" % (scene, name, out[scene][name]["ms_per_step"]))
        ref = out[scene].get("reference", {}).get("ms_per_step")
        for name, v in out[scene].items():
            if ref and name != "reference":
                v["speedup_vs_reference"] = ref / v["ms_per_step"]
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
