#include <ceed/ceed.h>
