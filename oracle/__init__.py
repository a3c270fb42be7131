"""CPU oracle for the pycsou proximal-solver hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``pycsou_b200/`` may import this package.
The only legitimate importers are ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- and there only as the
checker or the reported CPU baseline, never as the product path.

Contents
--------
``pylops_standin``  numpy/scipy restatement of the pylops v1.x operators that the reference's
                    ``linop/diff.py``, ``linop/conv.py`` and ``linop/sampling.py`` delegate to
                    (pylops is an un-vendored dependency, ``requirements.txt:4`` pins ``>= 1.9.2``).
``np_oracle``       stand-alone numpy restatement of the solver loop, the functionals and the
                    composition rules of the hot path; every function cites the reference
                    ``file:line`` it follows.  This is what travels to the GPU box.
``ref_shim``        imports the UNMODIFIED reference from ``/root/reference`` (build container only)
                    behind numpy-2 aliases and the pylops stand-in; used to pin ``np_oracle`` and to
                    generate ``tests/golden/*.npz``.  Never used at run time on the GPU box.

Parity pinning status: operator/prox level is pinned by the reference's own doctest literals
(transcribed in ``tests/test_oracle_doctests.py``); solver level is pinned by fixtures generated
by running the reference classes themselves (``tests/golden/gen_golden.py``), because the
reference has no solver tests.
"""
