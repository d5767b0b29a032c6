// C wrapper around the reference's own partitioned convolver (ear::dsp::block_convolver,
// /root/reference/visr_bear/submodules/libear/include/ear/dsp/block_convolver.hpp) so that the
// oracle's UPOLS core can be checked against reference code that actually compiles here.
// This file contains no reference source; the reference .cpp files are compiled in place.
#include "ear/dsp/block_convolver.hpp"
#include "ear/fft.hpp"

using namespace ear::dsp::block_convolver;

extern "C" void ear_block_convolve(const float *x, int len, const float *taps, int ntaps, int period, float *y)
{
  Context ctx(period, ear::get_fft_kiss<float>());
  Filter filter(ctx, ntaps, taps);
  BlockConvolver conv(ctx, filter);
  for (int b = 0; b + period <= len; b += period) conv.process(x + b, y + b);
}

// crossfade from taps_a to taps_b starting at block `switch_block` (libear semantics: input-side fade)
extern "C" void ear_block_convolve_xfade(const float *x, int len, const float *taps_a, const float *taps_b, int ntaps,
                                         int period, int switch_block, float *y)
{
  Context ctx(period, ear::get_fft_kiss<float>());
  Filter fa(ctx, ntaps, taps_a), fb(ctx, ntaps, taps_b);
  BlockConvolver conv(ctx, fa);
  for (int b = 0, k = 0; b + period <= len; b += period, k++) {
    if (k == switch_block) conv.crossfade_filter(fb);
    conv.process(x + b, y + b);
  }
}
