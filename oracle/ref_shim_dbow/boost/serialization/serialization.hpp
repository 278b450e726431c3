// Header shim, TEST INFRASTRUCTURE ONLY: BowVector.h / FeatureVector.h befriend boost::serialization::access and carry
// a serialize() member template that is never instantiated here (map save / load is out of the hot path).
#pragma once
namespace boost {
namespace serialization {
class access {};
template <class Base, class Derived> Base& base_object(Derived& d) { return d; }
}  // namespace serialization
}  // namespace boost
