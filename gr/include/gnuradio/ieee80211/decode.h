/* gr::ieee80211::decode -- public interface of the BCC decoder block.
 * Same factory and contract as the reference's include/gnuradio/ieee80211/decode.h:35
 * (make(bool ifdebug), one float input carrying demod's tags, message port "out";
 * lib/decode_impl.cc:60-520). */
#ifndef INCLUDED_IEEE80211_DECODE_H
#define INCLUDED_IEEE80211_DECODE_H

#include <gnuradio/block.h>
#include <gnuradio/ieee80211/api.h>

namespace gr {
namespace ieee80211 {

class IEEE80211_API decode : virtual public gr::block
{
public:
    typedef std::shared_ptr<decode> sptr;
    static sptr make(bool ifdebug);
};

} // namespace ieee80211
} // namespace gr

#endif
