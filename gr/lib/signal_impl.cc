/* signal_impl.cc -- legacy-signal block, 1 antenna(s): sync flags + raw samples in, packets in demod's input
 * format out (lib/signal_impl.cc:33-53 for the io signature and tag policy); the work is signal_core with nant = 1. */
#include "signal_impl.h"

#include <gnuradio/io_signature.h>

namespace gr {
namespace ieee80211 {

signal::sptr signal::make() { return gnuradio::make_block_sptr<signal_impl>(); }

signal_impl::signal_impl()
    : gr::block("signal", gr::io_signature::makev(2, 2, std::vector<int>{ sizeof(uint8_t), sizeof(gr_complex) }), gr::io_signature::make(1, 1, sizeof(gr_complex))),
      d_core(this, 1, "signal")
{
    set_tag_propagation_policy(block::TPP_DONT);
}

signal_impl::~signal_impl() {}

void signal_impl::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    d_core.forecast(noutput_items, ninput_items_required);
}

int signal_impl::general_work(int noutput_items,
                              gr_vector_int& ninput_items,
                              gr_vector_const_void_star& input_items,
                              gr_vector_void_star& output_items)
{
    return d_core.work(noutput_items, ninput_items, input_items, output_items);
}

} // namespace ieee80211
} // namespace gr
