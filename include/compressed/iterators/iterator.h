// compressed/iterators/iterator.h -- the reference keeps channel_iterator<T> here (iterators/iterator.h:28-330); in this
// tree it lives next to channel<T>, whose contexts and super-chunk it works on.
#pragma once
#include "../channel.h"
