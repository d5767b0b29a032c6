"""bear_cuda_process_device_static_scratch: the render call for hosts that may not allocate.

A plug-in host or sound-card thread calls the renderer from its audio callback, where a device
allocation could block past the period deadline; other embedders get their device memory once at
start-up.  For them the library offers a variant of bear_cuda_process_device whose only per-call
scratch (the partial output spectra) lives in an array built into the library, so the call
allocates nothing.  This test drives the variant through the C ABI next to the ordinary entry point
on identical inputs, two calls in a row (the second call finds the first call's leftovers in the
shared array and must not be affected by them), and checks the size limit.

Collected last on purpose: a run that stops at the first failure has then run everything else.
"""
import ctypes

import numpy as np
import pytest

from bear_b200 import control, data_file

pytestmark = pytest.mark.gpu


def _scene(T, B, seed):
    data = data_file.from_tree(data_file.synthetic_tree(seed=2, num_views=8, brir_length=1300))
    ctl = control.HostControl(data)
    rng = np.random.default_rng(seed)
    N, M, H = 6, 2, 4
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=[0, .5, 0, 1, 0, .3], ds_speakers=[2, 5],
                    hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=45.0 * np.arange(T))
    audio = np.concatenate([control.synth_audio(c, 2 * T * B, s) for c, s in ((N, 0), (M, 100), (H, 200))])
    return data, N, M, H, tb, audio


def test_static_scratch_variant_matches_process_device_twice_in_a_row():
    from bear_b200 import device_io
    from bear_b200.renderer import Renderer, STATIC_SCRATCH_BYTES
    T, B = 4, 512
    data, N, M, H, tb, audio = _scene(T, B, seed=11)
    with Renderer(data, N, M, H, B, max_batch_blocks=T) as plain, Renderer(data, N, M, H, B, max_batch_blocks=T) as noalloc:
        need = T * 2 * 2 * noalloc.info.cmac_splits * B * 8
        assert need <= STATIC_SCRATCH_BYTES
        io_a, io_b = device_io.DeviceBuffers(plain), device_io.DeviceBuffers(noalloc)
        for call in range(2):
            chunk = audio[:, call * T * B:(call + 1) * T * B]
            for io in (io_a, io_b):
                io.load_audio(chunk)
                io.load_tables(tb)
            plain.process_device(T)
            plain.synchronize()
            want = io_a.read_output(T)
            noalloc.process_device_static_scratch(T)
            noalloc.synchronize()
            got = io_b.read_output(T)
            assert np.abs(want).max() > 0.05
            # same kernels, same order of operations: the tolerance of the parity tests is generous here
            np.testing.assert_allclose(got, want, atol=1e-5, rtol=0)
        # and the renderer is still consistent with the ordinary entry point afterwards
        for io in (io_a, io_b):
            io.load_audio(audio[:, :T * B])
            io.load_tables(tb)
        plain.process_device(T)
        noalloc.process_device(T)
        np.testing.assert_allclose(io_b.read_output(T), io_a.read_output(T), atol=1e-5, rtol=0)


@pytest.mark.parametrize("chunk_blocks", [0, 5])
def test_static_scratch_variant_long_calls_match_the_oracle_and_allocate_nothing(oracle, chunk_blocks):
    """Calls of 12 periods take the convolution kernel that groups the periods by view (its work-item
    tables index into the built-in array with the number of splits the device picks); a chunked call
    reuses the array chunk after chunk.  Checked against the CPU oracle, not only against the other entry
    point, and the free device memory must not move across the calls.  (The host side of the call is
    allocation-free by construction -- no containers, error text in a buffer reserved at creation; the
    process heap cannot be asserted on from here because the Python call itself and the CUDA runtime's
    own bookkeeping allocate.)"""
    import torch
    from bear_b200 import device_io
    from bear_b200.renderer import Renderer, STATIC_SCRATCH_BYTES
    T, B = 12, 512
    data, N, M, H, tb, audio = _scene(T, B, seed=13)
    ref_r = oracle.RefRenderer(data, N, M, H, B)
    with Renderer(data, N, M, H, B, max_batch_blocks=T, cmac_splits=4, chunk_blocks=chunk_blocks) as r:
        assert T * 2 * 2 * r.info.cmac_splits * B * 8 <= STATIC_SCRATCH_BYTES
        io = device_io.DeviceBuffers(r)
        fn = r._l.bear_cuda_process_device_static_scratch
        for call in range(3):
            chunk = audio[:, (call % 2) * T * B:((call % 2) + 1) * T * B]
            io.load_audio(chunk)
            io.load_tables(tb)
            r.synchronize()
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            rc = fn(r._h, ctypes.c_uint32(T))          # the bare C call: nothing of the Python wrapper in between
            free1 = torch.cuda.mem_get_info()[0]
            assert rc == 0
            if call > 0:                               # (the very first launch of a kernel loads its module)
                assert free1 == free0, "device memory was allocated or freed by the call"
            r.synchronize()
            got = io.read_output(T)
            want = ref_r.render(chunk[:N], chunk[N:N + M], chunk[N + M:], tb.blocks())
            assert np.abs(want).max() > 0.05
            assert np.abs(got - want).max() <= 1e-5
        info = r.get_info()
        if chunk_blocks == 0:   # one launch of 12 periods: the grouped kernel ran
            assert info.cmac_items_last > 0 and info.cmac_splits_last <= r.info.cmac_splits
    ref_r.close()


def test_static_scratch_variant_rejects_oversize_calls():
    from bear_b200.renderer import BearCudaError, Renderer, STATIC_SCRATCH_BYTES
    T, B = 64, 2048
    data, N, M, H, tb, audio = _scene(2, B, seed=12)
    with Renderer(data, N, M, H, B, max_batch_blocks=T, cmac_splits=4) as r:
        assert T * 2 * 2 * r.info.cmac_splits * B * 8 > STATIC_SCRATCH_BYTES
        with pytest.raises(BearCudaError, match="static scratch too small"):
            r.process_device_static_scratch(T)
        r.process_device_static_scratch(2)   # within the limit: fine
        r.synchronize()
        with pytest.raises(ValueError, match="max_batch_blocks"):
            r.process_device_static_scratch(T + 1)
