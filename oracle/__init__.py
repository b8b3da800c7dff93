"""CPU oracle for the coarse-to-fine matching head -- TEST INFRASTRUCTURE ONLY.

Nothing in `featurematching_b200/` may import this package.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs use it, and there only as the checker / the timed CPU baseline, never as
the product path.

Contents
--------
matcher_oracle.py   independent restatement (torch-CPU, fp32 or fp64) of the three
                    reference modules on the hot path:
                      network/utils/coarse_matching_new.py:30-143
                      network/module/fine_preprocess.py:10-62
                      network/utils/fine_matching_new.py:13-79
reference_shims.py  imports the UNMODIFIED reference modules from /root/reference
                    (present only in the build container) behind import shims for
                    the third-party names that are absent there (matplotlib,
                    kornia).  Used by make_golden.py and by tests that pin the
                    restatement against the real reference when it is available.
make_golden.py      generates tests/golden/*.npz by running the real reference.
compare.py          parity rules (bit-exact indices up to documented ties,
                    1e-3 px coordinates, 1e-3 relative confidences).

Parity pinning
--------------
The reference ships no tests, golden vectors or fixtures, so
the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF, run in the build
container through reference_shims.py; those outputs are committed under
tests/golden/ together with the generating script (make_golden.py).  The one
piece of third-party arithmetic (kornia's `spatial_expectation2d` and
`create_meshgrid`, un-pinned version, fine_matching_new.py:63-66) is restated in
reference_shims.py from its published semantics.
"""
