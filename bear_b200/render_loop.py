"""The block loop of `bear-render` over planar numpy audio (bear/render_cli.py:398-465 restated).

The reference's BEARRenderer.render is called with one block of interleaved file samples: it pushes,
per rendering item, metadata blocks into the renderer until add_*_block answers False (the block
starts after the coming period: keep it for the next call), zero-pads a short last block, routes the
tracks and calls process().  ADM parsing, item selection and track routing are EAR's (not in this
repository); here an item is a list of already-converted `bear_b200.api` blocks and its channels are
rows of a planar float32 array.
"""
import numpy as np

from . import api


class _Pusher:
    """render_cli.py:405-420 push_single_channel for one kind of item."""

    def __init__(self, items, add_func, channels_per_item):
        self.items = [iter(blocks) for blocks in items]
        self.next = [None] * len(items)
        self.add = add_func
        self.nch = channels_per_item

    def push(self):
        start = 0
        for i, it in enumerate(self.items):
            while True:
                if self.next[i] is None:
                    self.next[i] = next(it, None)
                if self.next[i] is not None and self.add(i, self.next[i]):
                    self.next[i] = None
                    continue
                break
            start += self.nch[i]


class BlockRenderer:
    """objects_items / direct_speakers_items: one list of ObjectsInput / DirectSpeakersInput per channel;
    hoa_items: one list of HOAInput per stream, `hoa_channels[i]` channels each."""

    def __init__(self, data_path, objects_items=(), direct_speakers_items=(), hoa_items=(), hoa_channels=(),
                 block_size=512, sample_rate=48000, device=0):
        config = api.Config()
        config.num_objects_channels = len(objects_items)
        config.num_direct_speakers_channels = len(direct_speakers_items)
        config.num_hoa_channels = int(sum(hoa_channels))
        config.period_size = block_size
        config.sample_rate = sample_rate
        config.data_path = str(data_path)
        config.device = device
        self.config = config
        self.block_size = block_size
        self.renderer = api.Renderer(config)
        r = self.renderer
        self.pushers = [_Pusher(objects_items, r.add_objects_block, [1] * len(objects_items)),
                        _Pusher(direct_speakers_items, r.add_direct_speakers_block, [1] * len(direct_speakers_items)),
                        _Pusher(hoa_items, r.add_hoa_block, list(hoa_channels))]

    def render(self, objects, direct_speakers, hoa):
        """One block: [channels, <= block_size] float32 each (a short last block is zero-padded,
        render_cli.py:445-448).  Returns [2, block_size]."""
        for p in self.pushers:
            p.push()

        def pad(x, nch):
            x = np.zeros((nch, 0), np.float32) if x is None else np.asarray(x, np.float32)
            if x.shape[1] < self.block_size:
                full = np.zeros((nch, self.block_size), np.float32)
                full[:, :x.shape[1]] = x
                return full
            return np.ascontiguousarray(x)

        c = self.config
        output = np.zeros((2, self.block_size), np.float32)
        self.renderer.process(pad(objects, c.num_objects_channels), pad(direct_speakers, c.num_direct_speakers_channels),
                              pad(hoa, c.num_hoa_channels), output)
        return output

    def render_all(self, objects, direct_speakers, hoa, num_samples):
        """Whole planar signals, block by block; like the reference the tail after the last input
        block is dropped (render_cli.py:467-468 get_tail)."""
        B = self.block_size
        out = np.zeros((2, -(-num_samples // B) * B), np.float32)
        for s in range(0, num_samples, B):
            sl = lambda x: None if x is None else x[:, s:min(s + B, num_samples)]
            out[:, s:s + B] = self.render(sl(objects), sl(direct_speakers), sl(hoa))
        return out[:, :num_samples] if num_samples % B == 0 else out
