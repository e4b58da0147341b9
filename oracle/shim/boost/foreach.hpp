/* Minimal stand-in for <boost/foreach.hpp> so the reference's hot-path translation unit
 * (ObservationModelFB.cpp and the headers it pulls in) compiles without Boost, which is
 * not installed in this image.  Test infrastructure only (oracle/_ref build). */
#ifndef DINDEL_B200_SHIM_BOOST_FOREACH_HPP
#define DINDEL_B200_SHIM_BOOST_FOREACH_HPP
#define DINDEL_SHIM_CAT2(a, b) a##b
#define DINDEL_SHIM_CAT(a, b) DINDEL_SHIM_CAT2(a, b)
/* Single-statement range loop over an STL container that honours `break` and `continue`:
 * the outer loop walks the iterator, the inner one-shot loop binds the element; a `break`
 * inside the body leaves the flag set, which stops the outer loop as well. */
#define BOOST_FOREACH(decl, cont)                                                              \
    for (bool DINDEL_SHIM_CAT(_fe_brk_, __LINE__) = false; !DINDEL_SHIM_CAT(_fe_brk_, __LINE__);       \
         DINDEL_SHIM_CAT(_fe_brk_, __LINE__) = true)                                               \
        for (__typeof__((cont).begin()) DINDEL_SHIM_CAT(_fe_it_, __LINE__) = (cont).begin();       \
             !DINDEL_SHIM_CAT(_fe_brk_, __LINE__) && DINDEL_SHIM_CAT(_fe_it_, __LINE__) != (cont).end(); \
             ++DINDEL_SHIM_CAT(_fe_it_, __LINE__))                                                 \
            if ((DINDEL_SHIM_CAT(_fe_brk_, __LINE__) = true), false) {} else                       \
                for (decl = *DINDEL_SHIM_CAT(_fe_it_, __LINE__); DINDEL_SHIM_CAT(_fe_brk_, __LINE__);  \
                     DINDEL_SHIM_CAT(_fe_brk_, __LINE__) = false)
#endif
