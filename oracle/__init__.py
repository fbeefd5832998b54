"""CPU oracle for the DiffSep predictor-corrector sampler hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``ditsep_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker
or as the timed CPU baseline -- never as the product path.

The oracle is a functional fp32 PyTorch-CPU restatement of the reference's
algorithm (``/root/reference/src``), written against the reference's
``state_dict`` key layout so that the very same weights drive the reference,
the oracle and the CUDA path.  Every function cites the reference file:line it
follows.

Parity pin: the reference ships no tests, golden vectors or fixtures for this
path, so the pin is the reference itself, imported and
executed in the build container by ``oracle/pin_against_reference.py`` (which
also writes the committed fixtures under ``tests/golden/``).  On the GPU box
``/root/reference`` does not exist; there the oracle is checked against those
fixtures.
"""
