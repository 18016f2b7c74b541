/* signal2_impl.h -- gr::ieee80211::signal2 on the GPU batch library (2 receive antenna(s)). */
#ifndef INCLUDED_IEEE80211_SIGNAL2_IMPL_H
#define INCLUDED_IEEE80211_SIGNAL2_IMPL_H

#include <gnuradio/ieee80211/signal2.h>
#include "signal_core.h"

namespace gr {
namespace ieee80211 {

class signal2_impl : public signal2
{
public:
    signal2_impl();
    ~signal2_impl() override;

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
