// Library initialisation (reference: src/init.cpp:24-67).  The reference
// prints its notices and exit(1)s unless TENZING_ACK_NOTICE is set; as a
// library that may be imported from Python we print the notice once to stderr
// and continue.  Executables call init(argc, argv) first.
#pragma once

namespace tenzing {
void init();
void init(int &argc, char **&argv);
bool notice_acknowledged();
} // namespace tenzing
