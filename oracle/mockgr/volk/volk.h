/* Mock: the reference includes <volk/volk.h> but calls nothing from it on this path. */
#pragma once
