#include <glm/gtx/quaternion.hpp>
