/* gr::ieee80211::demod -- public interface of the SISO demodulator block.
 *
 * Same factory, base class and stream contract as the reference's
 * include/gnuradio/ieee80211/demod.h:35 (make(int mupos, int mugid), one gr_complex input, one float
 * output, tags in / tags out as lib/demod_impl.cc:59-342 reads and writes them); the implementation
 * behind it (gr/lib/demod_impl.cc) gathers whole packets and runs them on the GPU through
 * include/w80211rx.h instead of the per-symbol CPU loop. */
#ifndef INCLUDED_IEEE80211_DEMOD_H
#define INCLUDED_IEEE80211_DEMOD_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API demod : virtual public gr::block
{
public:
    typedef std::shared_ptr<demod> sptr;
    /* mupos / mugid: MU-MIMO user position and group id of the reference block.  mupos (0 or 1) selects this
     * receiver's stream of a VHT MU PPDU (lib/demod_impl.cc:347-379); mugid is kept for the
     * signature only, the reference never reads it either. */
    static sptr make(int mupos, int mugid);
};

} // namespace ieee80211
} // namespace gr

#endif
