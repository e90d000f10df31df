// placeholder, filled below
#pragma once
