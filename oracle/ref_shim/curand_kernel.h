// Empty stand-in: the reference's device sources are compiled as plain host C++ (see ref_host.cpp).
