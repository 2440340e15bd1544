// Test-only stand-in for the external JDQZPP eigensolver (not vendored by the reference, not on the hot path): the two
// members Continuation.H calls when an eigensolver has been set, which the tests never do.
#ifndef IEMIC_COMPAT_JDQZ_HPP
#define IEMIC_COMPAT_JDQZ_HPP
template <class Interface> class JDQZ {
 public:
  void printParameters() {}
  void solve() {}
};
#endif
