import logging

logger = logging.getLogger('chroma_b200')
