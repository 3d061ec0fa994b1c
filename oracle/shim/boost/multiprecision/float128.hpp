// TEST INFRASTRUCTURE ONLY (oracle). place.h:23 includes this header under GCC but every use of it is
// commented out (place.h:27-28); Boost is not installed in this image, so an empty stand-in suffices.
