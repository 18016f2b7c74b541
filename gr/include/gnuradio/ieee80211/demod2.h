/* gr::ieee80211::demod2 -- public interface of the two-antenna (2x2) demodulator block.
 * Same factory and stream contract as the reference's include/gnuradio/ieee80211/demod2.h:35
 * (make(), two gr_complex inputs, one float output; lib/demod2_impl.cc:58-349). */
#ifndef INCLUDED_IEEE80211_DEMOD2_H
#define INCLUDED_IEEE80211_DEMOD2_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API demod2 : virtual public gr::block
{
public:
    typedef std::shared_ptr<demod2> sptr;
    static sptr make();
};

} // namespace ieee80211
} // namespace gr

#endif
