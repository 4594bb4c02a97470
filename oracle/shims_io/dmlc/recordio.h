// Stand-in for <dmlc/recordio.h>: nothing of it is used by the text formats. Test infrastructure only.
#pragma once
