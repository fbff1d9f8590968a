#pragma once
#include <symengine/basic.h>
