// Test-infrastructure stub (NOT product code)
#include <pmt/pmt.h>
