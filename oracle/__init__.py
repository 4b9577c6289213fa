"""CPU oracle for the pinot GCN -> GP hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in ``pinot_b200`` imports this package.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs may import it, and only as the checker or
the timed CPU baseline -- never as the thing shipped.

Contents
--------
``dgl_ref``    restatement (plain torch, CPU) of the DGL pieces pinot delegates to.  DGL is an
               UNPINNED third-party dependency that is absent from /root/reference and cannot be
               installed here (requirements.txt:1, devtools/conda-envs/pinot.yaml:26; API usage
               brackets it to DGL 0.4.3 - 0.6.x).  Its documented layer math is restated; the
               reference holds no golden vectors for it, so this half is "parity unpinned" and is
               cross-checked against dense-adjacency algebra instead (tests/test_oracle.py).
``gp_ref``     restatement of pinot/regressors/{gaussian_process_regressor,kernels/*}.py in modern
               torch, fp32 or fp64.  PINNED: tests/golden/*.pt hold outputs of the reference's own
               code (imported from /root/reference by tests/golden/make_golden.py in the build
               container) and tests/test_oracle.py checks gp_ref against them.
``net_ref``    Sequential + Net composition of the two, loading the reference's state_dict keys.
``ref_loader`` imports the unmodified reference with stub modules (build container only).
"""
