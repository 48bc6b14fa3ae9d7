"""CPU oracle for the ModelOpt quantization hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(``tensorrt-model-optimizer_b200/``) may import this package.  It is used by
``tests/``, by ``__graft_entry__.smoke()`` and by the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the *checker* and the *timed CPU
baseline*, never as the thing shipped.

Every function restates one reference function in plain fp32 torch / numpy on
CPU and cites the reference ``file:line`` it follows (paths relative to the
reference checkout).  Parity status: **pinned** — ``oracle/gen_golden.py``
imports the real reference in the build container (CPU) and writes its outputs
to ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every oracle
function against those vectors bit-for-bit.  The one exception is the NVFP4
*dynamic* fake-quant, which has no CPU path in the reference
(``modelopt/torch/quantization/tensor_quant.py:172``): its oracle is an IEEE
restatement of the Triton kernel, pinned (a) structurally against the
reference's own E2M1 golden vectors and (b) bit-for-bit against outputs of the
reference's compiled Triton kernel captured on a B200
(``tests/golden/nvfp4_triton_b200.npz``, see ``oracle/gen_golden_gpu.py``).
"""
