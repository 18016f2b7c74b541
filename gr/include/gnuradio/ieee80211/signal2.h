/* gr::ieee80211::signal2 -- public interface of the legacy-signal block for two receive antennas.
 * Same factory and stream contract as the reference's include/gnuradio/ieee80211/signal2.h:35
 * (make(); input 0 = sync flags (uint8, tags rad/snr/rssi), inputs 1-2 = gr_complex samples of the two
 * antennas; two gr_complex outputs, tags on output 0; lib/signal2_impl.cc:62-222). */
#ifndef INCLUDED_IEEE80211_SIGNAL2_H
#define INCLUDED_IEEE80211_SIGNAL2_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API signal2 : virtual public gr::block
{
public:
    typedef std::shared_ptr<signal2> sptr;
    static sptr make();
};

} // namespace ieee80211
} // namespace gr

#endif
