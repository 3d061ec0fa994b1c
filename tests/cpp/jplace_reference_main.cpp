// TEST HARNESS: writes the shared test data with the REFERENCE's jplace writer
// (/root/reference/epik/src/epik/jplace.cpp, compiled where it lies, over the oracle's i2l shim and whatever
// RapidJSON the container has). Only used to (re)generate tests/golden/jplace_reference.jplace.
#include <fstream>
#include <epik/place.h>
#include <epik/jplace.h>
#include "jplace_case.h"

int main(int argc, char** argv)
{
    if (argc < 2) { return 2; }
    jplace_case::write_all<epik::io::jplace_writer, epik::impl::placed_collection, epik::impl::placed_sequence>(argv[1]);
    return 0;
}
