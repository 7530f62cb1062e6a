// TEST INFRASTRUCTURE: see NVX/nvx.h of this shim
#include <NVX/nvx.h>
