// TEST INFRASTRUCTURE.  mesh/LayeredMarchingCubesNoColor.cpp:46 only names boost::tuple in an unused typedef.
#ifndef GSDF_ORACLE_SHIM_BOOST_TUPLE_HPP_
#define GSDF_ORACLE_SHIM_BOOST_TUPLE_HPP_
namespace boost { template <class... Ts> struct tuple {}; }
#endif
