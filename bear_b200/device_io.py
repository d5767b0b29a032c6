"""Access to a renderer's device staging areas (bear_cuda_get_device_io) through torch tensors.

torch is plumbing here: it wraps the raw device pointers the C ABI hands out so that producers
that already live on the GPU (the bench's resident-input leg, the NCCL reduction of partial stereo
mixes) can read and write them without a host round trip."""
import numpy as np
import torch


def _wrap(ptr, shape, dtype, device):
    """torch view of foreign device memory via __cuda_array_interface__."""
    shape = tuple(int(s) for s in shape)
    if not ptr or int(np.prod(shape)) == 0:
        return torch.empty(shape, dtype=dtype, device=device)
    typestr = {torch.float32: "<f4", torch.int32: "<i4"}[dtype]

    class _Holder:
        __cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(_Holder(), device=device)


class DeviceBuffers:
    """Torch views of one renderer's staging areas, ordered on the renderer's own stream."""

    def __init__(self, renderer, num_speakers=24, num_hoa_decoder_channels=16, device_index=0):
        self.r = renderer
        io = renderer.device_io()
        dev = torch.device("cuda", device_index)
        self.device = dev
        T, N, M, H = renderer.max_batch_blocks, renderer.N, renderer.M, renderer.H
        L, A = num_speakers, num_hoa_decoder_channels
        self.stream = torch.cuda.ExternalStream(io.stream, device=dev)
        self.offset = io.audio_offset
        self.audio_in = _wrap(io.audio_in, (max(N + M + H, 1), io.audio_stride), torch.float32, dev)[:N + M + H]
        self.audio_out = _wrap(io.audio_out, (2, io.out_stride), torch.float32, dev)
        f, i = torch.float32, torch.int32
        self.tables = {
            "direct_gains": _wrap(io.direct_gains, (T, L, N), f, dev),
            "diffuse_gains": _wrap(io.diffuse_gains, (T, L, N), f, dev),
            "direct_delays": _wrap(io.direct_delays, (T, 2 * N), f, dev),
            "direct_speakers_gains": _wrap(io.direct_speakers_gains, (T, L, M), f, dev),
            "direct_speakers_delays": _wrap(io.direct_speakers_delays, (T, 2 * M), f, dev),
            "static_delays": _wrap(io.static_delays, (T, 2 * L), f, dev),
            "brir_index": _wrap(io.brir_index, (T,), i, dev),
            "hoa_gains": _wrap(io.hoa_gains, (T, A, H), f, dev),
        }

    def load_audio(self, audio):
        """audio: [N+M+H, T*B] (numpy or torch, host or device) -> staging area."""
        a = torch.as_tensor(audio)
        with torch.cuda.stream(self.stream):
            self.audio_in[:, self.offset:self.offset + a.shape[1]].copy_(a, non_blocking=True)

    def load_tables(self, tb):
        with torch.cuda.stream(self.stream):
            for k, v in tb.items():
                t = torch.as_tensor(v if isinstance(v, torch.Tensor) else np.ascontiguousarray(v))
                if t.numel():
                    self.tables[k][:t.shape[0]].copy_(t, non_blocking=True)

    def output(self, T):
        return self.audio_out[:, :T * self.r.B]

    def read_output(self, T, pipelined=False):
        """Copy the output of the last call to the host.  pipelined: the call was
        process_device_pipelined, whose output completes on the renderer's result stream."""
        st = self.stream
        if pipelined:
            st = torch.cuda.ExternalStream(self.r.result_stream(), device=self.device)
        with torch.cuda.stream(st):
            out = self.output(T).cpu()
        st.synchronize()
        return out.numpy()
