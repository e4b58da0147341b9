/* Minimal stand-in for <boost/tuple/tuple.hpp>: the reference only declares tuples on
 * paths outside the hot path.  Test infrastructure only (oracle/_ref build). */
#ifndef DINDEL_B200_SHIM_BOOST_TUPLE_HPP
#define DINDEL_B200_SHIM_BOOST_TUPLE_HPP
namespace boost {
template <class A, class B = int, class C = int, class D = int>
struct tuple {
    A a; B b; C c; D d;
    tuple() {}
    tuple(const A &a_, const B &b_) : a(a_), b(b_) {}
    tuple(const A &a_, const B &b_, const C &c_) : a(a_), b(b_), c(c_) {}
    tuple(const A &a_, const B &b_, const C &c_, const D &d_) : a(a_), b(b_), c(c_), d(d_) {}
};
}
#endif
