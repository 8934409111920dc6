"""Minimal stand-in for `ml_collections` (absent from this image, no network).

TEST INFRASTRUCTURE ONLY.  The reference's config files do
`ml_collections.ConfigDict()` and then plain attribute reads/writes; this
attribute-dict covers exactly that.
"""


class ConfigDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def __delattr__(self, k):
        del self[k]
