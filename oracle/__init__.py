"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the Hopfield association that
joejenb/Hop-VAE imports from the third-party package ``hflayers``.

Nothing in the shipped path (``hflayers/``, ``hop-vae_b200/``) may import this
package.  The only allowed importers are ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- there it
is the *checker* (or the timed CPU baseline), never the product.

Provenance / pinning
--------------------
The arithmetic of the hot path is NOT in ``/root/reference``: ``HopVAE.py:7``
does ``from hflayers import HopfieldLayer`` (ml-jku/hopfield-layers, pip name
``hopfield-layers``; un-vendored, un-pinned, not installable offline).  The
modules here restate that package's published algorithm for the options the
reference uses (``HopVAE.py:149-170``) plus the north-star options
(``scaling``, ``update_steps_max``, ``update_steps_eps``, non-static Q/K
projections).

Pins (see ``tests/test_oracle_pins.py``):
  * structural: ``checkpoints/MNIST-28.ckpt`` loads ``strict=True`` into the
    restated module tree (all 33 Hopfield keys, names and shapes);
  * behavioural: the trained-checkpoint known-answer test of SURVEY.md B.4
    (embedding_recon_loss 1.08e-5, reconstruction MSE 5.19e-4 on the
    deterministic stroke batch, and 10^2-10^3 x worse for any other beta/T);
  * algebraic: equality with ``torch.nn.functional.scaled_dot_product_attention``
    (an independent implementation that is on the box) and with the fp64 NumPy
    restatement in ``hopfield_f64.py``.
Options the reference does not exercise (T>0 stop criterion, Q/K projections)
are restated from knowledge of the public package only: for those, parity is
"unpinned" (against this restatement).
"""
