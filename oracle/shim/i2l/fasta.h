// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — place.cpp:11 includes <i2l/fasta.h> but uses nothing
// from it beyond seq_record. The real batch reader is part of the absent i2l (main.cpp:332).
#ifndef I2L_SHIM_FASTA_H
#define I2L_SHIM_FASTA_H
#include <i2l/seq_record.h>
#endif
