/* Mock of gr::fft::fft_complex_fwd (GNU Radio >= 3.10 wraps FFTW here; FFTW is
 * absent in this image).  The transform is the project's canonical 64-point
 * DFT (fft64_canon.h): unnormalised, natural order, the same mathematical
 * result FFTW returns, with a fixed FP32 rounding sequence.  Other sizes fall
 * back to a double-precision DFT rounded to float.  Test infrastructure only. */
#pragma once
#include <cmath>
#include <complex>
#include <vector>
#include <gnuradio/gr_complex.h>
#include "fft64_canon.h"

namespace gr {
namespace fft {
class fft_complex_fwd
{
    int d_n;
    std::vector<gr_complex> d_in, d_out;

public:
    fft_complex_fwd(int fft_size, int nthreads = 1) : d_n(fft_size), d_in(fft_size), d_out(fft_size)
    {
        (void)nthreads;
    }
    gr_complex* get_inbuf() { return d_in.data(); }
    gr_complex* get_outbuf() { return d_out.data(); }
    int inbuf_length() const { return d_n; }
    int outbuf_length() const { return d_n; }
    void execute()
    {
        if (d_n == 64) {
            c8_fft64(reinterpret_cast<const float*>(d_in.data()), reinterpret_cast<float*>(d_out.data()));
            return;
        }
        for (int k = 0; k < d_n; k++) {
            double sr = 0, si = 0;
            for (int n = 0; n < d_n; n++) {
                double a = -2.0 * M_PI * (double)n * (double)k / (double)d_n;
                sr += d_in[n].real() * cos(a) - d_in[n].imag() * sin(a);
                si += d_in[n].real() * sin(a) + d_in[n].imag() * cos(a);
            }
            d_out[k] = gr_complex((float)sr, (float)si);
        }
    }
};
} // namespace fft
} // namespace gr
