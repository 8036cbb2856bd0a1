"""Glue for running the UNMODIFIED reference training loop (``main.py``, ``SimpleDDPG``) and ``inference.py``
on gsc_b200: import stand-ins for the reference's un-installable dependencies (``compat/shims``), the module
that takes the place of ``src/rlsp/agents/models.py`` (``compat/reference_models.py``) and the runner
(``compat/run_reference.py``).  Nothing here is on the CUDA hot path and nothing in ``gsc_b200`` imports it."""
