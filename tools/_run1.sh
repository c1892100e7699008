mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_e2e.py::test_vae_encode_and_validate_step tests/test_gpu_pair_misc.py::test_im2col_s2_pad_after_gemm_is_vae_downsample -x -q -s > gpurun_out/enc.log 2>&1; echo "rc=$?"; grep -E "parity|passed|failed|Error|error" gpurun_out/enc.log | head -20
