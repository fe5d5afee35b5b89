// TEST INFRASTRUCTURE.  Intentionally empty (see tuple.hpp).
