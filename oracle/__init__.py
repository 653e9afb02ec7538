"""CPU oracle for the StyleGAN2-ADA synthesis hot path of VisCoIN-Next.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker or as the timed CPU baseline - never as the thing shipped.  The product
package (``viscoin-next_b200/``) must not import this package.

What is here
------------
* ``ref_ops.py``    - restatement (torch, CPU) of the reference's ``impl='ref'``
                      operator path: bias_act, upfirdn2d (+ filter2d / upsample2d
                      / downsample2d / setup_filter), conv2d_resample, fma,
                      modulated_conv2d.
* ``ref_net.py``    - functional restatement of the synthesis / mapping networks
                      that consume those ops, driven by a reference state_dict.
* ``kernel_math.py``- numpy restatement of the reference CUDA kernels' own
                      arithmetic (bias_act grad 0/1/2 table, upfirdn2d gather
                      index math) for small cases.
* ``make_golden.py``- imports the REAL reference from ``/root/reference`` (only
                      available in the build container) and writes the pinned
                      input/output vectors under ``tests/golden/``.

Parity status: PINNED.  The reference ships no tests / golden vectors of its
own, so the oracle is pinned against outputs of the
reference itself, generated in the build container by ``make_golden.py`` and
committed under ``tests/golden/``; ``tests/test_oracle_golden.py`` replays them.
Third-party arithmetic underneath both (ATen convolution) is the same torch
build on both sides and is not pinned by any reference test.
"""
