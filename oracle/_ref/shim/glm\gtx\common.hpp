#include <glm/gtx/common.hpp>
