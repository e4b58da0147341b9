/* Stand-in for <boost/math/special_functions/digamma.hpp> (Boost is not installed in this image;
 * the reference pins no Boost version: "parity unpinned" for this one function).
 * boost::math::digamma (call sites DInDel.cpp:2466,2472) forwards to the oracle's restatement of
 * the published algorithm (upward recurrence to x >= 10, then the asymptotic series),
 * oracle/dindel_oracle.c:ora_digamma, the same code the CUDA K3 mirrors.  Test infrastructure only. */
#ifndef DINDEL_B200_SHIM_BOOST_DIGAMMA_HPP
#define DINDEL_B200_SHIM_BOOST_DIGAMMA_HPP
extern "C" double ora_digamma(double x);
namespace boost { namespace math {
inline double digamma(double x) { return ora_digamma(x); }
}}
#endif
