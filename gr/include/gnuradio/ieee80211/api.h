/* Symbol visibility of the gnuradio-ieee80211 library: same macro name and meaning as the reference's
 * include/gnuradio/ieee80211/api.h:16-20, so the three block headers below can be swapped in. */
#ifndef INCLUDED_IEEE80211_API_H
#define INCLUDED_IEEE80211_API_H

#include <gnuradio/attributes.h>

#if defined(gnuradio_ieee80211_EXPORTS)
#define IEEE80211_API __GR_ATTR_EXPORT
#else
#define IEEE80211_API __GR_ATTR_IMPORT
#endif

#endif
