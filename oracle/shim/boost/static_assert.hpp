/* oracle/shim: Boost is not installed; the reference only needs BOOST_STATIC_ASSERT here. */
#ifndef BOOST_STATIC_ASSERT
#define BOOST_STATIC_ASSERT(x) static_assert(x, #x)
#endif
