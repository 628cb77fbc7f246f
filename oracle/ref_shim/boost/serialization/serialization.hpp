// CPU ORACLE support (test infrastructure, not product code).
// Stand-in for <boost/serialization/serialization.hpp>: Thirdparty/DBoW2/DBoW2/BowVector.h and FeatureVector.h of the reference
// declare a serialize() member template against it (BowVector.h:62-67); nothing on the path instantiates it, so the two names it
// mentions are all that is needed to compile those headers unmodified (boost is not in this image).
#pragma once
namespace boost {
namespace serialization {
class access {};
template <class Base, class Derived>
Base& base_object(Derived& d) { return static_cast<Base&>(d); }
}  // namespace serialization
}  // namespace boost
