#include <gnuradio/thread/thread.h>
