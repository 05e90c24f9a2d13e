"""CPU oracle for the ocrd_keraslm rating path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``ocrd_keraslm_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU
baseline legs do, and only as the checker / the timed CPU arm.

Parity status: the *host logic* (windowing, padding, state carry, beam search,
path selection) is pinned against the reference's own ``ocrd_keraslm/lib/rating.py``
executed in the build container (``tests/golden/make_golden.py`` drives it through
``oracle.keras_shim``); the *cell arithmetic* (Keras 2.3 / TF 1.15, third-party and
absent from /root/reference) is a restatement cross-checked against
``torch.nn.LSTM`` on CPU -- "parity unpinned" for the arithmetic, see DESIGN.md.
"""
