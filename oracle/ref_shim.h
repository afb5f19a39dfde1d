/* Force-included when compiling the reference's own C files in place
 * (oracle/Makefile).  The reference headers are not self-contained:
 * src/cortex/tk_cortex_main.h:112 uses tk_vision_object_t and :331 uses
 * tk_contextual_reasoner_t without including their definitions, and
 * src/vision/tk_image_preprocessor.c calls malloc without <stdlib.h>. */
#include <stdlib.h>
#include "vision/tk_vision_pipeline.h"
typedef struct tk_contextual_reasoner_s tk_contextual_reasoner_t;
