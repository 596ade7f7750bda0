"""Top-level ``__psvWave_cpp`` -- the module name the reference package imports its extension class from
(psvWave/__init__.py:6).  The compiled module lives inside the package (psvwave_b200/__psvWave_cpp*.so); this is its
alias for code that imports the top-level name.  No fallback: without the built CUDA extension the import fails."""
from psvwave_b200.__psvWave_cpp import *  # noqa: F401,F403
from psvwave_b200.__psvWave_cpp import fdModel  # noqa: F401
