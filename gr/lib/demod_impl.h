/* demod_impl.h -- gr::ieee80211::demod on the GPU batch library (replaces lib/demod_impl.h of the reference). */
#ifndef INCLUDED_IEEE80211_DEMOD_IMPL_H
#define INCLUDED_IEEE80211_DEMOD_IMPL_H

#include <gnuradio/ieee80211/demod.h>
#include "demod_core.h"

namespace gr {
namespace ieee80211 {

class demod_impl : public demod
{
public:
    demod_impl(int mupos, int mugid);
    ~demod_impl() override;

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items,
                     gr_vector_int& ninput_items,
                     gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override;

private:
    int d_mupos, d_mugid; /* mupos selects the MU channel estimate; the reference never reads mugid either (demod_impl.cc:38-39) */
    demod_core d_core;
};

} // namespace ieee80211
} // namespace gr

#endif
