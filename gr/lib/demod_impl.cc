/* demod_impl.cc -- SISO demodulator block: io signature, tag policy and scheduler contract of the
 * reference (lib/demod_impl.cc:34-57); the work itself is demod_core + w80211_demod_batch(). */
#include "demod_impl.h"

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

demod::sptr demod::make(int mupos, int mugid) { return gnuradio::make_block_sptr<demod_impl>(mupos, mugid); }

demod_impl::demod_impl(int mupos, int mugid)
    : gr::block("demod", gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(1, 1, sizeof(float))),
      d_mupos(mupos),
      d_mugid(mugid),
      d_core(this, 1, "demod", mupos)
{
    set_tag_propagation_policy(block::TPP_DONT);
}

demod_impl::~demod_impl() {}

void demod_impl::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    d_core.forecast(noutput_items, ninput_items_required);
}

int demod_impl::general_work(int noutput_items,
                             gr_vector_int& ninput_items,
                             gr_vector_const_void_star& input_items,
                             gr_vector_void_star& output_items)
{
    return d_core.work(noutput_items, ninput_items, input_items, output_items);
}

} // namespace ieee80211
} // namespace gr
