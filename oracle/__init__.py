"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the raygun hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl
reference`` legs may import it, and only as the checker or the reported CPU
baseline -- never as the thing measured or shipped.

Contents
--------
``restate.py``     a functional, CPU, plain-``torch.nn.functional`` restatement
                   of the reference's networks / losses / block geometry, driven
                   by a ``state_dict`` + a small spec dict.  Travels to the GPU
                   box (it has no dependency on ``/root/reference``).
``ref_loader.py``  imports the REFERENCE modules verbatim from
                   ``/root/reference/src`` (package ``__init__``s bypassed).
                   Works only in the build container; used to pin
                   ``restate.py`` and to generate ``tests/golden/*.npz``.
``make_golden.py`` the script that generated the committed golden vectors.
``synth.py``       synthetic EM-like uint8 volumes (SURVEY.md 8d recipe).

Parity status: PINNED by differential execution against the imported
reference modules (the reference's own tests hold no numeric goldens, only
structural constants -- SURVEY.md 8c -- and those are checked too).
"""
