/* gr::ieee80211::signal -- public interface of the legacy-signal block (one antenna).
 * Same factory and stream contract as the reference's include/gnuradio/ieee80211/signal.h:35
 * (make(); input 0 = sync flags (uint8, tags rad/snr/rssi), input 1 = gr_complex samples; one gr_complex
 * output; lib/signal_impl.cc:62-212). */
#ifndef INCLUDED_IEEE80211_SIGNAL_H
#define INCLUDED_IEEE80211_SIGNAL_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API signal : virtual public gr::block
{
public:
    typedef std::shared_ptr<signal> sptr;
    static sptr make();
};

} // namespace ieee80211
} // namespace gr

#endif
