/* demod2_impl.h -- gr::ieee80211::demod2 (two receive antennas) on the GPU batch library. */
#ifndef INCLUDED_IEEE80211_DEMOD2_IMPL_H
#define INCLUDED_IEEE80211_DEMOD2_IMPL_H

#include <gnuradio/ieee80211/demod2.h>
#include "demod_core.h"

namespace gr {
namespace ieee80211 {

class demod2_impl : public demod2
{
public:
    demod2_impl();
    ~demod2_impl() override;

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override;

private:
    demod_core d_core;
};

} // namespace ieee80211
} // namespace gr

#endif
