"""CPU oracle for the GP-EI candidate-scoring hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, in numpy/scipy, the reference algorithm for the path
``Space.sample -> transform -> normalise -> GP posterior -> acquisition -> argmax``
(SURVEY.md section 8a).  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product (``metaopt_b200``) never does and has no CPU fallback.

Parity status
-------------
* sampling / transform rows (a1-a11): pinned against the reference's own
  known-answer tests and against outputs of the unmodified reference imported
  from ``/root/reference`` (``oracle/make_golden.py`` -> ``tests/golden/*.npz``).
* GP fit / posterior / acquisition rows (a13-a17): the arithmetic lives in
  third-party packages that are NOT in ``/root/reference`` and are not pinned by it
  (``orion.algo.skopt`` -> ``scikit-optimize`` -> ``scikit-learn``; the reference's
  ``setup.py:51`` pins none).  The reference holds no test or golden vector for it:
  **parity unpinned** by the reference.  The oracle is anchored instead on the
  on-box scikit-learn 1.9.0 ``GaussianProcessRegressor`` (the real code the
  reference's plugin would run) plus a restatement of skopt's published
  ``gaussian_ei / gaussian_pi / gaussian_lcb`` formulas.
"""
