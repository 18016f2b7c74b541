/* iqconv.cu -- sc16 -> fc32 conversion of a batch (16-bit IQ on the wire).
 *
 * The host feed is PCIe-bound: 8 bytes per complex sample in fc32, 2.15 bytes per decoded bit at C4.  USRPs
 * deliver sc16 natively (UHD converts to fc32 on the host for the reference's flowgraph, examples/rx.grc:744),
 * so the batch may arrive as int16 pairs plus one scale per packet and is expanded here, in HBM:
 * sample = ((float)re * scale, (float)im * scale) -- int16 -> binary32 is exact and the multiply is one rounded
 * binary32 operation, the same a host would do, so the reference fed those products sees identical input.
 * One block per packet; 4 bytes read and 8 written per sample, coalesced. */
#include "common.cuh"

__global__ void __launch_bounds__(256) sc16_to_fc32_kernel(int npkt, int nant, const w80211_pkt_in* __restrict__ pkts,
                                                           const short2* __restrict__ raw0, const short2* __restrict__ raw1,
                                                           float2* __restrict__ out0, float2* __restrict__ out1,
                                                           unsigned long long nsamples)
{
    const int p = blockIdx.x;
    if (p >= npkt) return;
    const w80211_pkt_in pk = pkts[p];
    if (pk.nsamp <= 0 || pk.iq_offset + (unsigned long long)pk.nsamp > nsamples) return; /* the header stage flags it */
    const float sc = pk.iq_scale;
    for (int a = 0; a < nant; a++) {
        const short2* src = (a ? raw1 : raw0) + pk.iq_offset;
        float2* dst = (a ? out1 : out0) + pk.iq_offset;
        for (int i = threadIdx.x; i < pk.nsamp; i += blockDim.x) {
            const short2 v = __ldg(src + i);
            dst[i] = make_float2(__fmul_rn((float)v.x, sc), __fmul_rn((float)v.y, sc));
        }
    }
}

void w_launch_sc16_to_fc32(cudaStream_t st, int npkt, int nant, const w80211_pkt_in* pkts, const void* raw0, const void* raw1,
                           float2* out0, float2* out1, unsigned long long nsamples)
{
    if (npkt <= 0) return;
    sc16_to_fc32_kernel<<<npkt, 256, 0, st>>>(npkt, nant, pkts, (const short2*)raw0, (const short2*)raw1, out0, out1, nsamples);
}
