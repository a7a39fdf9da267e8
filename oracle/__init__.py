"""CPU oracle for the emotional-vits latent->waveform path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``vits_b200/`` may import this
package; it is used by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` as the checker
and the timed CPU baseline, never as the product path.
"""
