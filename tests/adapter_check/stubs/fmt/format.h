#pragma once
#include <fmt/core.h>
