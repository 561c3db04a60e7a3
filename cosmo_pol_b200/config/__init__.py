from .cfg import init, sanity_check, CONFIG, DEFAULTS, VALID_VALUES
