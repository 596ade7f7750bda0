"""``import psvWave`` for scripts and tests written against the reference package (psvWave/__init__.py:1-125 of
larsgeb/psvWave): the same names -- ``fdModel``, ``get_dictionary``, ``from_dictionary``, ``write_dictionary`` -- served by
the B200 solver of ``psvwave_b200``.  With the repository root on ``sys.path`` (and as the working directory, for the
relative ``tests/test_configurations/...ini`` paths they use) the reference's own ``tests/python/test_copy.py`` and
``test_descent.py`` run unmodified; ``tests/test_gpu_reference_suite.py`` does exactly that where the reference checkout
is present.  The matplotlib plot helpers (psvWave/__init__.py:128-334) are out of scope for this path and are not here.
"""
from psvwave_b200 import (__version__, fdModel, from_dictionary, get_dictionary,  # noqa: F401
                          write_dictionary)
