"""TEST INFRASTRUCTURE — the CPU oracle of the ORB front-end path (see oracle/orb_oracle.hpp).
Never imported by the product package."""
