/* ref_driver.cc -- C API around the UNMODIFIED reference blocks.
 *
 * Built by oracle/Makefile together with
 *   /root/reference/lib/{cloud80211phy,signal_impl,signal2_impl,demod_impl,
 *                        demod2_impl,decode_impl}.cc
 * (compiled where they lie, against oracle/mockgr) into oracle/_ref/libref80211.so.
 * Nothing here restates reference arithmetic: it only creates blocks through
 * their public make() factories and plays scheduler (mock_scheduler.h).
 * Test infrastructure / CPU baseline only -- never linked into the product.
 *
 * With -DW80211_DRIVER_OURS the same stream / block / scheduler entry points are compiled against
 * THIS repo's drop-in blocks (gr/include, gr/lib) instead, into oracle/_ours/libours80211.so: the
 * block-level equivalence tests then feed identical streams to both libraries (tests/test_gr_blocks.py).
 */
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/decode.h>
#include <gnuradio/ieee80211/demod.h>
#include <gnuradio/ieee80211/demod2.h>
#include <gnuradio/ieee80211/signal.h>
#include <gnuradio/ieee80211/signal2.h>
#ifdef W80211_DRIVER_OURS
#include <gnuradio/ieee80211/detect.h>
#endif
#ifndef W80211_DRIVER_OURS
#include <gnuradio/ieee80211/sync.h>
#include <gnuradio/ieee80211/trigger.h>
#include "cloud80211phy.h"
#endif
#include "mock_scheduler.h"

using gr::mock_stream;

struct ref_block {
    std::shared_ptr<gr::block> blk;
    std::vector<std::vector<uint8_t>> msgs;
    bool honor_forecast = false; /* trigger / sync index their input by noutput_items */
};

static std::string g_key_tmp;

extern "C" {

/* ------------------------------------------------------------ streams */
mock_stream* ref_stream_new(int itemsize, long nitems, const void* data)
{
    mock_stream* s = new mock_stream(itemsize);
    if (nitems > 0) {
        s->data.resize((size_t)nitems * itemsize);
        if (data)
            memcpy(s->data.data(), data, s->data.size());
    }
    return s;
}
void ref_stream_free(mock_stream* s) { delete s; }
long ref_stream_nitems(const mock_stream* s) { return (long)s->nitems(); }
int ref_stream_itemsize(const mock_stream* s) { return s->itemsize; }
const void* ref_stream_data(const mock_stream* s) { return s->data.data(); }

static void add_tag(mock_stream* s, long offset, const char* key, pmt::pmt_t v)
{
    gr::tag_t t;
    t.offset = (uint64_t)offset;
    t.key = pmt::mp(key);
    t.value = v;
    s->tags.push_back(t);
}
void ref_stream_add_tag_long(mock_stream* s, long offset, const char* key, long v)
{
    add_tag(s, offset, key, pmt::from_long(v));
}
void ref_stream_add_tag_float(mock_stream* s, long offset, const char* key, float v)
{
    add_tag(s, offset, key, pmt::from_float(v));
}
void ref_stream_add_tag_c32(mock_stream* s, long offset, const char* key, const float* data, int n)
{
    add_tag(s, offset, key, pmt::init_c32vector((size_t)n, reinterpret_cast<const std::complex<float>*>(data)));
}
/* tags were added in ascending offset order (or are put in that order now): lookups become binary searches */
void ref_stream_sort_tags(mock_stream* s)
{
    std::stable_sort(s->tags.begin(), s->tags.end(), [](const gr::tag_t& a, const gr::tag_t& b) { return a.offset < b.offset; });
    s->tags_sorted = true;
}
int ref_stream_ntags(const mock_stream* s) { return (int)s->tags.size(); }
long ref_stream_tag_offset(const mock_stream* s, int i) { return (long)s->tags[i].offset; }
const char* ref_stream_tag_key(const mock_stream* s, int i)
{
    g_key_tmp = pmt::symbol_to_string(s->tags[i].key);
    return g_key_tmp.c_str();
}
/* 0 long, 1 double, 2 c32 vector, -1 other */
int ref_stream_tag_kind(const mock_stream* s, int i)
{
    switch (s->tags[i].value->kind) {
    case pmt::K_LONG: return 0;
    case pmt::K_DOUBLE: return 1;
    case pmt::K_C32VEC: return 2;
    default: return -1;
    }
}
long ref_stream_tag_long(const mock_stream* s, int i) { return s->tags[i].value->l; }
double ref_stream_tag_double(const mock_stream* s, int i) { return s->tags[i].value->d; }
int ref_stream_tag_c32(const mock_stream* s, int i, float* out, int maxn)
{
    const auto& v = s->tags[i].value->c32;
    int n = (int)std::min<size_t>(v.size(), (size_t)maxn);
    if (out && n > 0) memcpy(out, v.data(), (size_t)n * 8);
    return (int)v.size();
}

/* ------------------------------------------------------------- blocks */
ref_block* ref_block_new(const char* kind, int a0, int a1)
{
    ref_block* b = new ref_block;
    std::string k(kind);
    if (k == "signal")
        b->blk = gr::ieee80211::signal::make();
    else if (k == "signal2")
        b->blk = gr::ieee80211::signal2::make();
#ifndef W80211_DRIVER_OURS
    else if (k == "trigger" || k == "sync") {
        if (k == "trigger")
            b->blk = gr::ieee80211::trigger::make();
        else
            b->blk = gr::ieee80211::sync::make();
        b->honor_forecast = true;
    }
#endif
#ifdef W80211_DRIVER_OURS
    else if (k == "detect")
        b->blk = gr::ieee80211::detect::make();
#endif
    else if (k == "demod")
        b->blk = gr::ieee80211::demod::make(a0, a1);
    else if (k == "demod2")
        b->blk = gr::ieee80211::demod2::make();
    else if (k == "decode")
        b->blk = gr::ieee80211::decode::make(a0 != 0);
    else {
        delete b;
        return nullptr;
    }
    return b;
}
void ref_block_free(ref_block* b) { delete b; }

/* Run the block over whole streams; nout = 0 for a sink (decode).
 * Returns the number of general_work calls, or -1 on an exception. */
long ref_block_run(ref_block* b, mock_stream** ins, int nin, mock_stream** outs, int nout, long chunk_in, long chunk_out)
{
    std::vector<const mock_stream*> vin(ins, ins + nin);
    std::vector<mock_stream*> vout;
    for (int o = 0; o < nout; o++) vout.push_back(outs[o]);
    long calls;
    try {
        calls = gr::mock_run(*b->blk, vin, vout, chunk_in, chunk_out, b->honor_forecast);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_block_run: exception: %s\n", e.what());
        return -1;
    }
    for (auto& m : b->blk->mock_msgs) {
        /* every message on this path is pmt::cons(dict, blob) */
        pmt::pmt_t payload = pmt::cdr(m.second);
        b->msgs.push_back(payload->blob);
    }
    b->blk->mock_msgs.clear();
    return calls;
}
/* demod -> decode joined by a bounded edge, run in turn by one thread (mock_run_chain) or each on a thread of its own
 * (mock_run_chain_threads): the streaming forms the block-path benchmark and the chain tests use.  Messages land in `down`. */
long ref_chain_run(ref_block* up, ref_block* down, mock_stream** ins, int nin, long ring_items, long chunk_in, long chunk_out, int threads)
{
    std::vector<const mock_stream*> vin(ins, ins + nin);
    long calls;
    try {
        calls = threads ? gr::mock_run_chain_threads(*up->blk, *down->blk, vin, ring_items, chunk_in, chunk_out)
                        : gr::mock_run_chain(*up->blk, *down->blk, vin, ring_items, chunk_in, chunk_out);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_chain_run: exception: %s\n", e.what());
        return -1;
    }
    for (auto& m : down->blk->mock_msgs) down->msgs.push_back(pmt::cdr(m.second)->blob);
    down->blk->mock_msgs.clear();
    return calls;
}
int ref_block_nmsgs(const ref_block* b) { return (int)b->msgs.size(); }
int ref_block_msg_len(const ref_block* b, int i) { return (int)b->msgs[i].size(); }
const uint8_t* ref_block_msg_data(const ref_block* b, int i) { return b->msgs[i].data(); }
void ref_block_clear_msgs(ref_block* b) { b->msgs.clear(); }

#ifndef W80211_DRIVER_OURS
/* ------------------------------------------- PHY library entry points
 * Thin pass-throughs to free functions of the reference's cloud80211phy.cc,
 * for unit-level parity tests of the restatement and of the CUDA kernels. */
void ref_fill_mod_legacy(int mcs, int len, c8p_mod* m) { signalParserL(mcs, len, m); }

/* fields of c8p_mod that matter on the path, in a fixed order */
static void mod_to_ints(const c8p_mod& m, int* o)
{
    o[0] = m.format; o[1] = m.sumu; o[2] = m.ampdu; o[3] = m.nSym; o[4] = m.nSymSamp;
    o[5] = m.nSD; o[6] = m.nSS; o[7] = m.nLTF; o[8] = m.mcs; o[9] = m.len;
    o[10] = m.mod; o[11] = m.cr; o[12] = m.nBPSCS; o[13] = m.nDBPS; o[14] = m.nCBPS; o[15] = m.nCBPSS;
}
void ref_parse_legacy(int mcs, int len, int* out16)
{
    c8p_mod m;
    memset(&m, 0, sizeof(m));
    signalParserL(mcs, len, &m);
    mod_to_ints(m, out16);
}
int ref_parse_ht(const uint8_t* bits48, int* out16)
{
    c8p_mod m;
    c8p_sigHt s;
    memset(&m, 0, sizeof(m));
    uint8_t b[48];
    memcpy(b, bits48, 48);
    if (!signalCheckHt(b)) return 0;
    signalParserHt(b, &m, &s);
    mod_to_ints(m, out16);
    return 1;
}
int ref_parse_vht(const uint8_t* sigA48, const uint8_t* sigB26, int* out16)
{
    c8p_mod m;
    c8p_sigVhtA s;
    memset(&m, 0, sizeof(m));
    uint8_t a[48], bb[26];
    memcpy(a, sigA48, 48);
    memcpy(bb, sigB26, 26);
    if (!signalCheckVhtA(a)) return 0;
    signalParserVhtA(a, &m, &s);
    signalParserVhtB(bb, &m);
    mod_to_ints(m, out16);
    return 1;
}
void ref_qam_to_llr(int mod, int nSD, const float* qam_in, float* llr_out)
{
    c8p_mod m;
    memset(&m, 0, sizeof(m));
    m.mod = mod;
    m.nSD = nSD;
    std::vector<gr_complex> q(nSD);
    memcpy(q.data(), qam_in, (size_t)nSD * 8);
    procSymQamToLlr(q.data(), llr_out, &m);
}
/* which: 0 legacy, 1 non-legacy stream 1, 2 non-legacy stream 2 */
void ref_deint(int which, int nCBPSS, const float* in, float* out)
{
    c8p_mod m;
    memset(&m, 0, sizeof(m));
    m.nCBPS = nCBPSS;
    m.nCBPSS = nCBPSS;
    std::vector<float> tmp(in, in + nCBPSS);
    if (which == 0)
        procSymDeintL2(tmp.data(), out, &m);
    else if (which == 1)
        procSymDeintNL2SS1(tmp.data(), out, &m);
    else
        procSymDeintNL2SS2(tmp.data(), out, &m);
}
void ref_depas2(int nBPSCS, int nCBPSS, const float* in0, const float* in1, float* out)
{
    static float in[C8P_MAX_N_SS][C8P_MAX_N_CBPSS];
    c8p_mod m;
    memset(&m, 0, sizeof(m));
    m.nBPSCS = nBPSCS;
    m.nCBPSS = nCBPSS;
    memcpy(in[0], in0, (size_t)nCBPSS * 4);
    memcpy(in[1], in1, (size_t)nCBPSS * 4);
    procSymDepasNL(in, out, &m);
}
void ref_sv_decode(const float* llr, uint8_t* bits, int trellis)
{
    std::vector<float> tmp(llr, llr + 2 * (size_t)trellis);
    SV_Decode_Sig(tmp.data(), bits, trellis);
}
void ref_bcc_encode(const uint8_t* in, uint8_t* out, int len)
{
    std::vector<uint8_t> tmp(in, in + len);
    bccEncoder(tmp.data(), out, len);
}
void ref_sv_tables(int* next128, int* output128)
{
    for (int i = 0; i < 64; i++)
        for (int b = 0; b < 2; b++) {
            next128[2 * i + b] = SV_STATE_NEXT[i][b];
            output128[2 * i + b] = SV_STATE_OUTPUT[i][b];
        }
}
void ref_pilot_p(float* out127) { memcpy(out127, PILOT_P, sizeof(float) * 127); }
/* Read-only views of the reference's constant tables, for tests/test_tables.py.
 * Returns the number of entries written (as doubles), 0 for an unknown name. */
int ref_table(const char* name, double* out, int maxn)
{
    std::string k(name);
    const int* ip = nullptr;
    const float* fp = nullptr;
    int n = 0;
    if (k == "mapDeintLegacyBpsk") { ip = mapDeintLegacyBpsk; n = 48; }
    else if (k == "mapDeintLegacyQpsk") { ip = mapDeintLegacyQpsk; n = 96; }
    else if (k == "mapDeintLegacy16Qam") { ip = mapDeintLegacy16Qam; n = 192; }
    else if (k == "mapDeintLegacy64Qam") { ip = mapDeintLegacy64Qam; n = 288; }
    else if (k == "mapDeintNonlegacyBpsk") { ip = mapDeintNonlegacyBpsk; n = 52; }
    else if (k == "mapDeintNonlegacyQpsk") { ip = mapDeintNonlegacyQpsk; n = 104; }
    else if (k == "mapDeintNonlegacy16Qam") { ip = mapDeintNonlegacy16Qam; n = 208; }
    else if (k == "mapDeintNonlegacy64Qam") { ip = mapDeintNonlegacy64Qam; n = 312; }
    else if (k == "mapDeintNonlegacy256Qam") { ip = mapDeintNonlegacy256Qam; n = 416; }
    else if (k == "mapDeintVhtSigB20") { ip = mapDeintVhtSigB20; n = 52; }
    else if (k == "LTF_NL_28_F_FLOAT") { fp = LTF_NL_28_F_FLOAT; n = 64; }
    else if (k == "LTF_NL_28_F_FLOAT2") { fp = LTF_NL_28_F_FLOAT2; n = 64; }
    else if (k == "LTF_L_26_F_FLOAT") { fp = LTF_L_26_F_FLOAT; n = 64; }
    else if (k == "PILOT_P") { fp = PILOT_P; n = 127; }
    if (n > maxn) n = maxn;
    for (int i = 0; i < n; i++) out[i] = ip ? (double)ip[i] : (double)fp[i];
    return n;
}
void ref_intel_vhtb20(const uint8_t* in52, uint8_t* out52)
{
    uint8_t tmp[52];
    memcpy(tmp, in52, 52);
    procIntelVhtB20(tmp, out52);
}
/* procNLSigDemodDeint (cloud80211phy.cc:629-648) on caller-supplied FFT bins */
void ref_nlsig_demod(const float* sym1, const float* sym2, const float* h, float* llrht96, float* llrvht96)
{
    std::vector<gr_complex> s1(64), s2(64), hh(64);
    memcpy(s1.data(), sym1, 512);
    memcpy(s2.data(), sym2, 512);
    memcpy(hh.data(), h, 512);
    procNLSigDemodDeint(s1.data(), s2.data(), hh, llrht96, llrvht96);
}
int ref_check_crc8(const uint8_t* bits42)
{
    uint8_t tmp[42];
    memcpy(tmp, bits42, 42);
    return checkBitCrc8(tmp, 34, &tmp[34]) ? 1 : 0;
}

#endif /* !W80211_DRIVER_OURS */

} /* extern "C" */
