"""Scene runner: `python -m mantaflow_b200 scene.py [args...]` runs a Mantaflow scene file AS-IS on the device path,
the way the reference's executable does (`manta scene.py`, source/pwrapper/pymain.cpp:61-112: argv[0] = the scene file,
the file is executed as `__main__` with the `manta` module importable, GUI = False in a build without Qt).
"""
import os
import sys


def run_scene(path, argv=None, globs=None):
    """Executes the bytes of `path` unchanged.  `from manta import *` resolves to mantaflow_b200.manta."""
    from . import manta as M
    import types
    mod = sys.modules.get("manta")
    if mod is None or not getattr(mod, "__file__", "").startswith(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))):
        mod = types.ModuleType("manta")
        mod.__dict__.update({k: getattr(M, k) for k in dir(M) if not k.startswith("_")})
        mod.__all__ = [k for k in dir(M) if not k.startswith("_")]
        sys.modules["manta"] = mod
    with open(path, "rb") as f:
        src = f.read()
    g = {"__name__": "__main__", "__file__": path, "__builtins__": __builtins__}
    if globs:
        g.update(globs)
    old_argv = sys.argv
    sys.argv = [path] + list(argv or [])
    sys.path.insert(0, os.path.dirname(os.path.abspath(path)))
    try:
        exec(compile(src, path, "exec"), g)
    finally:
        sys.argv = old_argv
    return g


def main():
    if len(sys.argv) < 2:
        sys.stderr.write("Usage: python -m mantaflow_b200 <scene.py> [args]\n")
        return 1
    run_scene(sys.argv[1], sys.argv[2:])
    return 0


if __name__ == "__main__":
    sys.exit(main())
