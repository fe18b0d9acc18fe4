#include <mitsuba/shim.h>
