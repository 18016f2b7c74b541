/* signal_impl.h -- gr::ieee80211::signal on the GPU batch library (1 receive antenna(s)). */
#ifndef INCLUDED_IEEE80211_SIGNAL_IMPL_H
#define INCLUDED_IEEE80211_SIGNAL_IMPL_H

#include <gnuradio/ieee80211/signal.h>
#include "signal_core.h"

namespace gr {
namespace ieee80211 {

class signal_impl : public signal
{
public:
    signal_impl();
    ~signal_impl() override;

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override;

private:
    signal_core d_core;
};

} // namespace ieee80211
} // namespace gr

#endif
