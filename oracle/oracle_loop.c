/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * Placeholder translation unit: the loop-closure pieces (discriminators, loop filters, lock detectors)
 * are checked directly against the reference's own objects in oracle/_ref/libgnssref.so. */
#include "oracle.h"
int orc_version(void) { return 1; }
