// TEST HARNESS: writes the shared test data with epik_b200's jplace writer (no GPU involved).
#include <epik/place.h>
#include <epik/jplace.h>
#include "jplace_case.h"

int main(int argc, char** argv)
{
    if (argc < 2) { return 2; }
    jplace_case::write_all<epik::io::jplace_writer, epik::impl::placed_collection, epik::impl::placed_sequence>(argv[1]);
    return 0;
}
