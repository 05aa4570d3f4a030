#include <pmt/pmt.h>
