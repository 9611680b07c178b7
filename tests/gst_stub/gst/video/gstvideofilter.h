#include <gst/gst.h>
