/* iqconv.cu -- sc16 -> fc32 conversion of a batch (16-bit IQ on the wire).
 *
 * The host feed is PCIe-bound: 8 bytes per complex sample in fc32, 2.15 bytes per decoded bit at C4.  USRPs
 * deliver sc16 natively (UHD converts to fc32 on the host for the reference's flowgraph, examples/rx.grc:744),
 * so the batch may arrive as int16 pairs plus one scale per packet and is expanded here, in HBM:
 * sample = ((float)re * scale, (float)im * scale) -- int16 -> binary32 is exact and the multiply is one rounded
 * binary32 operation, the same a host would do, so the reference fed those products sees identical input.
 * One block per packet; 4 bytes read and 8 written per sample, coalesced. */
#include <algorithm>
#include "common.cuh"

template <class RAW>
__global__ void __launch_bounds__(256) int_to_fc32_kernel(int npkt, int nant, const w80211_pkt_in* __restrict__ pkts,
                                                          const RAW* __restrict__ raw0, const RAW* __restrict__ raw1,
                                                          float2* __restrict__ out0, float2* __restrict__ out1,
                                                          unsigned long long nsamples)
{
    const int p = blockIdx.x;
    if (p >= npkt) return;
    const w80211_pkt_in pk = pkts[p];
    if (pk.nsamp <= 0 || pk.iq_offset + (unsigned long long)pk.nsamp > nsamples) return; /* the header stage flags it */
    const float sc = pk.iq_scale;
    for (int a = 0; a < nant; a++) {
        const RAW* src = (a ? raw1 : raw0) + pk.iq_offset;
        float2* dst = (a ? out1 : out0) + pk.iq_offset;
        for (int i = threadIdx.x; i < pk.nsamp; i += blockDim.x) {
            const RAW v = __ldg(src + i);
            dst[i] = make_float2(__fmul_rn((float)v.x, sc), __fmul_rn((float)v.y, sc));
        }
    }
}

/* bytes_per_sample: 4 = sc16, 2 = sc8 */
void w_launch_sc16_to_fc32(cudaStream_t st, int npkt, int nant, const w80211_pkt_in* pkts, const void* raw0, const void* raw1,
                           float2* out0, float2* out1, unsigned long long nsamples, int bytes_per_sample)
{
    if (npkt <= 0) return;
    if (bytes_per_sample == 2)
        int_to_fc32_kernel<char2><<<npkt, 256, 0, st>>>(npkt, nant, pkts, (const char2*)raw0, (const char2*)raw1, out0, out1, nsamples);
    else
        int_to_fc32_kernel<short2><<<npkt, 256, 0, st>>>(npkt, nant, pkts, (const short2*)raw0, (const short2*)raw1, out0, out1, nsamples);
}

/* ------------------------------------------------------------------------------------------------
 * S_COPY of the reference's signal / signal2 blocks (lib/signal_impl.cc:164-191, signal2_impl.cc:167-199):
 * sample k of an accepted packet = raw[sync + 224 + k] * (cosf(t), sinf(t)), t = (float)(k + 224) * rad,
 * on every antenna; plus the packet's "chan" tag copied from the per-flag table.  16 bytes of HBM traffic per
 * sample and antenna; blockIdx.y walks the packet in chunks so that long and short packets share one grid. */
#define CFO_CHUNK 2048
__global__ void __launch_bounds__(256) cfo_copy_kernel(int npkt, int nant, const w_sig_acc* __restrict__ acc,
                                                       const float2* __restrict__ in0, const float2* __restrict__ in1,
                                                       float2* __restrict__ out0, float2* __restrict__ out1,
                                                       const float2* __restrict__ chan_all, float2* __restrict__ chan_out)
{
    const int p = blockIdx.x;
    if (p >= npkt) return;
    const w_sig_acc a = acc[p];
    const int k0 = blockIdx.y * CFO_CHUNK;
    if (blockIdx.y == 0 && threadIdx.x < 64) chan_out[(size_t)p * 64 + threadIdx.x] = chan_all[(size_t)a.sync * 64 + threadIdx.x];
    if (k0 >= a.nsamp) return;
    const int k1 = min(a.nsamp, k0 + CFO_CHUNK);
    for (int k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
        float sn, cs;
        sincosf(__fmul_rn((float)(k + 224), a.rad), &sn, &cs);
        const cf32 w = c_make(cs, sn);
        const float2 v = __ldg(in0 + a.src + 224 + k);
        const cf32 r = c_mul(c_make(v.x, v.y), w);
        out0[a.dst + k] = make_float2(r.x, r.y);
        if (nant == 2) {
            const float2 v1 = __ldg(in1 + a.src + 224 + k);
            const cf32 r1 = c_mul(c_make(v1.x, v1.y), w);
            out1[a.dst + k] = make_float2(r1.x, r1.y);
        }
    }
}

void w_launch_cfo_copy(cudaStream_t st, int npkt, int nant, int max_nsamp, const w_sig_acc* acc, const float2* in0,
                       const float2* in1, float2* out0, float2* out1, const float2* chan_all, float2* chan_out)
{
    if (npkt <= 0) return;
    dim3 grid(npkt, max(1, (max_nsamp + CFO_CHUNK - 1) / CFO_CHUNK));
    cfo_copy_kernel<<<grid, 256, 0, st>>>(npkt, nant, acc, in0, in1, out0, out1, chan_all, chan_out);
}

/* Small host -> device uploads (the host plan of a batch) done by a kernel that reads the pinned host buffer directly:
 * on the copy engine they would queue behind the gigabytes of input samples that later batches have already enqueued. */
__global__ void __launch_bounds__(256) upload_words_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, size_t nwords)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

void w_launch_upload(cudaStream_t st, void* dst, const void* pinned_src, size_t bytes)
{
    if (bytes == 0) return;
    const size_t nwords = (bytes + 3) / 4; /* every plan array is a whole number of 32-bit words in buffers padded beyond that */
    int blocks = (int)std::min<size_t>((nwords + 255) / 256, 1184);
    upload_words_kernel<<<blocks, 256, 0, st>>>((uint32_t*)dst, (const uint32_t*)pinned_src, nwords);
}
