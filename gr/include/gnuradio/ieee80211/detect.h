/* gr::ieee80211::detect -- packet detection on the GPU as ONE block.
 *
 * Not a block of the reference: it takes the place of three of them plus a hierarchy --
 *     presiso (examples/presiso.grc)  ->  trigger (lib/trigger_impl.cc:60-117)  ->  sync (lib/sync_impl.cc:61-196)
 * -- because the per-sample streams between those (preac, preconj, the trigger's flag bytes) are what a GPU stage
 * removes.  Input: the received gr_complex samples.  Output: the sync block's output stream, item for item: one uint8 per
 * sample, 0x01 at "L-LTF start + 16" of every detected packet with the tags rad / snr / rssi (pmt::from_float) on that
 * item (sync_impl.cc:121-134), zero elsewhere.  In rx.grc it is wired as  source -> detect -> signal input 0, with the
 * source also on signal's input 1; signal / demod / decode stay as they are. */
#ifndef INCLUDED_IEEE80211_DETECT_H
#define INCLUDED_IEEE80211_DETECT_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API detect : virtual public gr::block
{
public:
    typedef std::shared_ptr<detect> sptr;
    static sptr make();
};

} // namespace ieee80211
} // namespace gr

#endif
