/* @file Platform.hpp

    Headless platform: no window; the event loop draws a fixed number of frames
    (reference include/Platform.hpp:15-22, platform/Linux.cpp:356-367).
    SPDX-License-Identifier: WTFPL
*/
#ifndef PLATFORM_HPP
#define PLATFORM_HPP

#include "Common.hpp"

VkResult PlatformCreateWindow(void);
VkResult PlatformEnterEventLoop(void);

#endif
