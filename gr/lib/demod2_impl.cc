/* demod2_impl.cc -- 2x2 demodulator block: two gr_complex inputs, one float output, tags read on
 * input 0 only (lib/demod2_impl.cc:33-57,73-77); the work is demod_core with nant = 2. */
#include "demod2_impl.h"

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

demod2::sptr demod2::make() { return gnuradio::make_block_sptr<demod2_impl>(); }

demod2_impl::demod2_impl()
    : gr::block("demod2", gr::io_signature::make(2, 2, sizeof(gr_complex)), gr::io_signature::make(1, 1, sizeof(float))),
      d_core(this, 2, "demod2")
{
    set_tag_propagation_policy(block::TPP_DONT);
}

demod2_impl::~demod2_impl() {}

void demod2_impl::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    d_core.forecast(noutput_items, ninput_items_required);
}

int demod2_impl::general_work(int noutput_items,
                              gr_vector_int& ninput_items,
                              gr_vector_const_void_star& input_items,
                              gr_vector_void_star& output_items)
{
    return d_core.work(noutput_items, ninput_items, input_items, output_items);
}

} // namespace ieee80211
} // namespace gr
