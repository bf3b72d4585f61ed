"""Import alias: ``import b2v`` -> the package in
``decentralised-verification-and-distributed-execution-of-large-scale-video-diffusion-models_b200/``
(whose directory name is not a valid Python identifier)."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_NAME = "decentralised-verification-and-distributed-execution-of-large-scale-video-diffusion-models_b200"
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module(PKG_NAME)
sys.modules[__name__] = _pkg
sys.modules.setdefault("b2v", _pkg)
# make ``import b2v.ops`` / ``from b2v.strategies import chunk_only`` resolve to the SAME module objects
for _sub in ("_lib", "ops", "checkpoint", "vae", "validator", "scheduler", "unet", "pipeline", "planner", "proof", "strategies", "strategies.common",
             "strategies.chunk_only", "strategies.fsdp_chunked_coherent", "strategies.fsdp_chunked",
             "strategies.fsdp", "build", "exchange", "miner"):
    try:
        _m = importlib.import_module(PKG_NAME + "." + _sub)
    except ModuleNotFoundError as _e:  # a submodule that does not exist (yet) is not an error of the alias
        if _e.name and _e.name.startswith(PKG_NAME):
            continue
        raise
    sys.modules["b2v." + _sub] = _m
