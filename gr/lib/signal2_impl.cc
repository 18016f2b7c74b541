/* signal2_impl.cc -- legacy-signal block, 2 antenna(s): sync flags + raw samples in, packets in demod's input
 * format out (lib/signal2_impl.cc:33-53 for the io signature and tag policy); the work is signal_core with nant = 2. */
#include "signal2_impl.h"

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

signal2::sptr signal2::make() { return gnuradio::make_block_sptr<signal2_impl>(); }

signal2_impl::signal2_impl()
    : gr::block("signal2", gr::io_signature::makev(3, 3, std::vector<int>{ sizeof(uint8_t), sizeof(gr_complex), sizeof(gr_complex) }), gr::io_signature::make(2, 2, sizeof(gr_complex))),
      d_core(this, 2, "signal2")
{
    set_tag_propagation_policy(block::TPP_DONT);
}

signal2_impl::~signal2_impl() {}

void signal2_impl::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    d_core.forecast(noutput_items, ninput_items_required);
}

int signal2_impl::general_work(int noutput_items,
                               gr_vector_int& ninput_items,
                               gr_vector_const_void_star& input_items,
                               gr_vector_void_star& output_items)
{
    return d_core.work(noutput_items, ninput_items, input_items, output_items);
}

} // namespace ieee80211
} // namespace gr
